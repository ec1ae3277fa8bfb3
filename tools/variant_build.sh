#!/bin/bash
# Development: gpurun_in_<name>.so = the library with rt_render.cu compiled with extra flags (rt_build.o / rt_abi.o of the last full build).
#   tools/variant_build.sh name -DRT_DONATE_FIRST=256 ...
set -e
name=$1; shift
here=$(cd "$(dirname "$0")/.." && pwd)
pk=$here/unity-raytracer_b200
/usr/local/cuda/bin/nvcc "$@" -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true \
  -Xcompiler -fPIC,-ffp-contract=off,-O2 -I $here/include -c $pk/csrc/rt_render.cu -o /tmp/var_$name.o
/usr/local/cuda/bin/nvcc -shared -o $here/gpurun_in_$name.so $pk/build/rt_build.o /tmp/var_$name.o $pk/build/rt_abi.o -gencode arch=compute_100a,code=sm_100a
echo built gpurun_in_$name.so
