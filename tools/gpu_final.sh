set -x
python -m pytest tests -m gpu -x -q > gpurun_out/final_gputest.log 2>&1; tail -3 gpurun_out/final_gputest.log
python bench.py > gpurun_out/final_bench_cfg3_n1.json 2> gpurun_out/final_bench_cfg3_n1.err; tail -c 300 gpurun_out/final_bench_cfg3_n1.err
python bench.py --workload cfg4i --steps 5 --warmup 3 > gpurun_out/final_bench_cfg4i_n1.json 2> gpurun_out/final_bench_cfg4i_n1.err; tail -c 300 gpurun_out/final_bench_cfg4i_n1.err
python bench.py --workload cfg5 --steps 3 --warmup 3 --no-commit > gpurun_out/final_bench_cfg5_n1.json 2>/dev/null
python bench.py --workload cfg2 --steps 20 --warmup 3 --no-commit > gpurun_out/final_bench_cfg2_n1.json 2>/dev/null
python bench.py --workload cfg4 --builder lbvh --steps 5 --warmup 3 --no-commit > gpurun_out/final_bench_cfg4_lbvh_n1.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final_launches_cfg3.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/final_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_trace --launch-skip 16 -c 4 -f -o gpurun_out/final_trace_cfg3 python bench.py --steps 2 --warmup 3 --no-cpu --no-commit > gpurun_out/final_ncu_full.log 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_reference.json 2>/dev/null
python - <<'PY'
import json
for f in ["cfg3","cfg4i","cfg5","cfg2","cfg4_lbvh"]:
    try:
        d=json.load(open(f"gpurun_out/final_bench_{f}_n1.json")); print(f, round(d["value"]), round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],3), d["roofline"].get("ms_per_step_by_kernel"))
    except Exception as e: print(f, "ERR", e)
d=json.load(open("gpurun_out/final_bench_reference.json")); print("reference", d["value"], d["cpu_baseline"]["cores"])
PY
