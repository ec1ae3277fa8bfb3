set -x
python bench.py > gpurun_out/r2b_bench_cfg3_n1.json 2> gpurun_out/r2b_bench_cfg3_n1.err; tail -c 300 gpurun_out/r2b_bench_cfg3_n1.err
python - <<'PY'
import json; d=json.load(open('gpurun_out/r2b_bench_cfg3_n1.json')); print('cfg3', d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['ms_per_step_by_kernel'], d.get('moving_camera',{}).get('value'))
PY
python bench.py --workload cfg4i --steps 5 --warmup 3 > gpurun_out/r2b_bench_cfg4i_n1.json 2> gpurun_out/r2b_bench_cfg4i_n1.err; tail -c 600 gpurun_out/r2b_bench_cfg4i_n1.err; head -c 1500 gpurun_out/r2b_bench_cfg4i_n1.json
python bench.py --workload cfg4i --builder lbvh --steps 5 --warmup 3 --no-cpu > gpurun_out/r2b_bench_cfg4i_lbvh_n1.json 2> gpurun_out/r2b_bench_cfg4i_lbvh_n1.err; tail -c 300 gpurun_out/r2b_bench_cfg4i_lbvh_n1.err
python bench.py --workload cfg5 --steps 3 --warmup 3 --no-commit > gpurun_out/r2b_bench_cfg5_n1.json 2> gpurun_out/r2b_bench_cfg5_n1.err; tail -c 300 gpurun_out/r2b_bench_cfg5_n1.err; head -c 800 gpurun_out/r2b_bench_cfg5_n1.json
python bench.py --workload cfg2 --steps 20 --warmup 3 --no-commit > gpurun_out/r2b_bench_cfg2_n1.json 2> gpurun_out/r2b_bench_cfg2_n1.err; tail -c 300 gpurun_out/r2b_bench_cfg2_n1.err
python bench.py --workload cfg4 --builder lbvh --steps 5 --warmup 3 --no-commit > gpurun_out/r2b_bench_cfg4_lbvh_n1.json 2> gpurun_out/r2b_bench_cfg4_lbvh_n1.err; tail -c 300 gpurun_out/r2b_bench_cfg4_lbvh_n1.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_cfg3.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-commit > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_trace --launch-skip 16 -c 4 -f -o gpurun_out/r2_trace_cfg3 python bench.py --steps 2 --warmup 3 --no-cpu --no-commit > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out/
