set -x
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2_bench_cfg3_n$N.json 2> gpurun_out/r2_bench_cfg3_n$N.err; tail -c 600 gpurun_out/r2_bench_cfg3_n$N.err
python - <<PY
import json; d=json.load(open('gpurun_out/r2_bench_cfg3_n$N.json'))
print('N$N value', d['value'], 'ms', d['ms_per_step'], 'by rank', d['ms_per_step_by_rank'])
print('e2e', d['e2e']['value'], d['e2e']['ms_per_step'])
x=d.get('extra',{})
print('cfg4', x.get('cfg4',{}).get('value'), x.get('cfg4',{}).get('ms_per_step'), x.get('cfg4',{}).get('e2e'))
print('in_process', x.get('in_process',{}).get('bit_identical'), x.get('in_process',{}).get('e2e'))
print('kernels', d['roofline'].get('ms_per_step_by_kernel'))
PY
