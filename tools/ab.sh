#!/bin/bash
# Development aid: A/B of alternative builds of the library on the GPU box (gpurun -- bash tools/ab.sh cfg3 lib1.so lib2.so ...)
cfg=$1; shift
for lib in "$@"; do
  for extra in "" "part=8 rank=3"; do
    echo "=== $lib $cfg $extra"
    python tools/gpu_probe.py $cfg 6 lib=$lib flush=1 timing=0 $extra 2>&1 | grep '"ms"' | tail -4 | cut -c1-90
  done
done
