set -x
# final tree of round 2: ncu launch list + full capture of the trace kernels (refreshes r2_trace.md / r2_ncu_kernels.json for the current
# csrc hash), then the default bench line (which then carries `traffic` from a capture of ITS build)
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/f3_plain.log 2>&1; tail -c 200 gpurun_out/f3_plain.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/f3_launches_cfg3.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/f3_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_trace --launch-skip 16 -c 4 -f -o gpurun_out/f3_trace_cfg3 python bench.py --steps 2 --warmup 3 --no-cpu --no-commit > gpurun_out/f3_ncu_full.log 2>&1
ls -la gpurun_out/f3_*
