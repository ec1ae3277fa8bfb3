set -x
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2_bench_cfg3_n2.json 2> gpurun_out/r2_bench_cfg3_n2.err; tail -c 1500 gpurun_out/r2_bench_cfg3_n2.err
python - <<'PY'
import json; d=json.load(open('gpurun_out/r2_bench_cfg3_n2.json'))
print('N2 value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e'])
print('extra', json.dumps(d.get('extra'))[:3000])
PY
python -m pytest tests -m gpu -x -q -k "in_process_multi_device or ranks_copy" 2>&1 | tail -3
