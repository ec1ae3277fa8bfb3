#!/bin/bash
# The GPU parity suite on the self-checking build (-DRT_CHECKED: every node / slot / tile / queue / pixel / instance index the render
# kernels form is compared with its bound where it is used; rt_core.h).  compute-sanitizer is closed on this GPU pool.
#   here:        RTB200_EXTRA_NVCC_FLAGS=-DRT_CHECKED RTB200_OUT=$PWD/gpurun_in_checked.so python unity-raytracer_b200/build.py --force
#   on the box:  bash tools/checked_run.sh > gpurun_out/r2_checked_build.txt 2>&1
export RTB200_LIB=${RTB200_LIB:-$PWD/gpurun_in_checked.so}
python -m pytest tests -m gpu -q -x -k "not c_smoke_test" 2>&1 | tail -6
