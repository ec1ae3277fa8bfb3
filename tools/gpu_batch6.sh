set -x
python tools/gpu_probe.py cfg4i 3 flush=1 builder=1 timing=1 2>&1 | tail -1
for lib in c10 c12 pk0; do python tools/gpu_probe.py cfg4i 3 flush=1 builder=1 timing=1 lib=/root/repo/gpurun_in_$lib.so 2>&1 | tail -1; done
python tools/gpu_probe.py cfg4i 3 flush=1 builder=1 timing=1 fuse_shade=0 2>&1 | tail -1
python tools/gpu_probe.py cfg4i 3 flush=1 builder=1 timing=0 2>&1 | tail -1
VC_BUILDERS=1 python tools/visit_counts.py /root/repo/gpurun_in_cnt.so cfg4i gpurun_out/r2_visit_counts_cfg4i.json 2>&1 | tail -2
python -m pytest tests/test_instancing.py -m gpu -x -q 2>&1 | tail -2
