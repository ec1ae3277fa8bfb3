set -x
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sorted_mirror or cfg5_reduced or deep_mirror or lane_count" 2>&1 | tail -3
for sq in 0 1; do python tools/gpu_probe.py cfg5 3 flush=1 builder=1 sort_queue=$sq timing=0 2>&1 | tail -3; done
for sq in 0 1; do python tools/gpu_probe.py cfg5 2 flush=1 builder=1 sort_queue=$sq timing=1 2>&1 | tail -1; done
VC_BUILDERS=1 python tools/visit_counts.py /root/repo/gpurun_in_cnt.so cfg4i gpurun_out/r2_visit_counts_cfg4i.json 2>&1 | tail -2
VC_BUILDERS=1 python tools/visit_counts.py /root/repo/gpurun_in_cnt.so cfg4ib gpurun_out/r2_visit_counts_cfg4ib.json 2>&1 | tail -2
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,sass__inst_executed_local_loads,sass__inst_executed_local_stores,sass__inst_executed_global_loads,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,sm__warps_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,launch__registers_per_thread
for w in cfg4i cfg4ib; do
  ncu --metrics $M --clock-control none -k regex:k_trace -s 4 -c 4 --csv --log-file gpurun_out/r2_ncu_ab_$w.csv python tools/gpu_probe.py $w 2 flush=1 timing=1 builder=1 > gpurun_out/ncu_ab_$w.log 2>&1
done
ncu --set full --clock-control none --import-source on -k regex:k_trace_shadow -s 2 -c 1 -f -o gpurun_out/r2_inst_shadow python tools/gpu_probe.py cfg4i 2 flush=1 timing=1 builder=1 > gpurun_out/ncu_inst_full.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_trace_queue -s 40 -c 2 -f -o gpurun_out/r2_queue_cfg5 python tools/gpu_probe.py cfg5 1 flush=1 timing=0 builder=1 sort_queue=0 > gpurun_out/ncu_queue_full.log 2>&1
bash tools/sanitize.sh > gpurun_out/r2_sanitizer.txt 2>&1; tail -5 gpurun_out/r2_sanitizer.txt
ls -la gpurun_out | tail -12
