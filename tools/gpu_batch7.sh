set -x
python tools/gpu_probe.py cfg3 5 flush=1 builder=1 timing=1 2>&1 | tail -2
for lib in g ni; do python tools/gpu_probe.py cfg3 5 flush=1 builder=1 timing=1 lib=/root/repo/gpurun_in_$lib.so 2>&1 | tail -2; done
python tools/gpu_probe.py cfg3 5 flush=1 builder=1 timing=0 2>&1 | tail -2
for lib in g ni; do python tools/gpu_probe.py cfg3 5 flush=1 builder=1 timing=0 lib=/root/repo/gpurun_in_$lib.so 2>&1 | tail -2; done
for lib in g ni; do python tools/gpu_probe.py cfg3 5 flush=1 builder=1 timing=0 part=8 rank=3 lib=/root/repo/gpurun_in_$lib.so 2>&1 | tail -1; done
python tools/gpu_probe.py cfg3 5 flush=1 builder=1 timing=0 part=8 rank=3 2>&1 | tail -1
