set -x
for lib in oct0 oct1 pk pkoct1; do python tools/gpu_probe.py cfg4i 3 flush=1 builder=1 timing=1 lib=/root/repo/gpurun_in_$lib.so 2>&1 | tail -1; done
python tools/gpu_probe.py cfg4i 3 flush=1 builder=1 timing=1 fuse_shade=0 2>&1 | tail -1
python tools/gpu_probe.py cfg4i 3 flush=1 builder=1 timing=1 ctas_per_sm=6 2>&1 | tail -1
python tools/gpu_probe.py cfg4i 3 flush=1 builder=1 timing=1 ctas_per_sm=4 2>&1 | tail -1
python tools/gpu_probe.py cfg4i 3 flush=1 builder=1 timing=1 packet=0 2>&1 | tail -1
python tools/gpu_probe.py cfg4ib 3 flush=1 builder=1 timing=1 2>&1 | tail -1
