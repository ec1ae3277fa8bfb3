#!/bin/bash
# AddressSanitizer + UndefinedBehaviorSanitizer over the host-side C++ that can run without a GPU: the oracle
# (oracle/rt_oracle.cpp) and the product's host/device headers as the CPU emulation instantiates them (tests/emu/rt_emu.cpp:
# rt_core.h traversal and shading, rt_build_steps.h LBVH steps, rt_sah.h SAH builder + pre-splitting, rt_instances.h BLAS / TLAS
# packing, rt_host_scene.h).  Sanitized copies are built under /tmp and loaded through RT_ORACLE_LIB / RT_EMU_LIB:
#     bash tools/sanitize_host.sh > profiles/r2_sanitizer_host.txt 2>&1
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=/tmp/rt_san; mkdir -p $OUT
SAN="-fsanitize=address,undefined -fno-sanitize-recover=undefined -fno-omit-frame-pointer -g -O1"
COMMON="-std=c++17 -fPIC -shared -pthread -ffp-contract=off -fno-fast-math"
g++ $COMMON $SAN -o $OUT/librt_oracle.so $ROOT/oracle/rt_oracle.cpp
g++ $COMMON $SAN -Wno-unused-function -o $OUT/librt_emu.so $ROOT/tests/emu/rt_emu.cpp
echo "built: $(ls $OUT)"
cd $ROOT
# (Python itself is not instrumented: leak detection off, everything else fatal)
LD_PRELOAD="$(gcc -print-file-name=libasan.so) $(gcc -print-file-name=libubsan.so)" ASAN_OPTIONS=detect_leaks=0:abort_on_error=0:halt_on_error=1 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1 \
  RT_ORACLE_LIB=$OUT/librt_oracle.so RT_EMU_LIB=$OUT/librt_emu.so \
  python -m pytest tests/test_oracle.py tests/test_emu.py tests/test_independent.py tests/test_instancing.py -x -q -m "not gpu" -p no:cacheprovider 2>&1 | tail -15
