set -x
python bench.py --workload cfg4i --steps 5 --warmup 3 > gpurun_out/final_bench_cfg4i_n1.json 2> gpurun_out/final_bench_cfg4i_n1.err; tail -c 300 gpurun_out/final_bench_cfg4i_n1.err
ncu --set full --clock-control none --import-source on -k regex:k_trace --launch-skip 16 -c 4 -f -o gpurun_out/final_trace_cfg3 python bench.py --steps 2 --warmup 3 --no-cpu --no-commit > gpurun_out/final_ncu_full.log 2>&1
python bench.py > gpurun_out/final_bench_cfg3_n1.json 2> gpurun_out/final_bench_cfg3_n1.err; tail -c 300 gpurun_out/final_bench_cfg3_n1.err
