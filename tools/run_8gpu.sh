mkdir -p gpurun_out
N=${1:-8}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 5 2>/dev/null | grep '^{' > gpurun_out/bench_${N}gpu.json
python - <<PY
import json
d = json.load(open('gpurun_out/bench_${N}gpu.json'))
print('N', d['n_gpus'], 'value', round(d['value']/1e6, 2), 'M evals/s', 'ms', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']), 'evals/s', round(d['e2e']['ms_per_step'], 2), 'ms', 'd2h/gpu', round(d['e2e']['d2h_gbs_per_gpu'], 1), 'GB/s', d['e2e']['host_placement'])
PY
lscpu | grep -i "^CPU(s)\|numa node(s)\|socket"
