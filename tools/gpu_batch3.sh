set -x
python -m pytest tests/test_instancing.py -m gpu -x -q 2>&1 | tail -3
python tools/gpu_probe.py cfg4i 4 flush=1 builder=1 timing=0 2>&1 | tail -2
python tools/gpu_probe.py cfg4i 3 flush=1 builder=1 timing=1 2>&1 | tail -1
python tools/gpu_probe.py cfg4i 3 flush=1 builder=0 timing=1 2>&1 | tail -1
for sq in 0 1 2 3; do python tools/gpu_probe.py cfg5 3 flush=1 builder=1 sort_queue=$sq timing=0 2>&1 | tail -1; done
for sq in 0 1 3; do python tools/gpu_probe.py cfg5 2 flush=1 builder=1 sort_queue=$sq timing=1 2>&1 | tail -1; done
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sorted_mirror or cfg5_reduced or deep_mirror" 2>&1 | tail -2
