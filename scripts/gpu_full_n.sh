# usage: bash scripts/gpu_full_n.sh N tag — the driver's command at N GPUs (C5 leg included)
mkdir -p gpurun_out
N=$1; TAG=$2
if [ "$N" = "1" ]; then
  python bench.py --gpus 1 --steps 200 --warmup 10 > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err
else
  timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 200 --warmup 10 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
fi
tail -c 600 gpurun_out/${TAG}_bench_n$N.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/${TAG}_bench_n$N.json') if l.startswith('{')][-1])
print('N', d['n_gpus'], 'value', round(d['value']/1e6,1), 'us', round(d['ms_per_step']*1e3,2), 'e2e', round(d['e2e']['value']/1e6,1), d['exchange'], 'mc', d.get('exchange_multicast'), 'wp', d.get('exchange_w_private'))
print('graph', d.get('graph_phase_ms_per_step'))
e=d['eval']; print('eval', round(e['value']/1e6,2), 'e2e', round(e['e2e']['value']/1e6,2), e['sharding'], e['cpu_baseline'].get('ranks_outside_tie_band'))
c=d['c5']; print('c5', round(c['value']/1e6,1), 'ms', round(c['ms_per_step'],3), 'e2e', round(c['e2e']['value']/1e6,1), 'frac', round(c['roofline']['frac'],3), c['phase_ms_per_step'], 'parity', c['parity'], 'eval', round(c['eval']['value']), c['eval']['sharding'])
PY
