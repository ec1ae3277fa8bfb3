#!/bin/bash
# A/B ncu counters of k_trace_shadow for two builds
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,sass__inst_executed_local_loads,sass__inst_executed_local_stores,sass__inst_executed_global_loads,sass__inst_executed_shared_loads,sass__inst_executed_shared_stores,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_imc_miss_per_issue_active.ratio,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed_op_branch.sum,lts__t_sectors.sum
for lib in "$@"; do
  python tools/gpu_probe.py cfg3 3 lib=/root/repo/$lib flush=1 timing=1 builder=1 > gpurun_out/plain_$(basename $lib).log 2>&1 || exit 1
  ncu --metrics $M --clock-control none -k regex:k_trace_shadow -s 4 -c 2 --csv --log-file gpurun_out/ncu_$(basename $lib).csv python tools/gpu_probe.py cfg3 3 lib=/root/repo/$lib flush=1 timing=1 builder=1 > gpurun_out/ncu_$(basename $lib).log 2>&1
done
