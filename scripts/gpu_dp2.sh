# usage: bash scripts/gpu_dp2.sh N [tag] — real-NVLink checks + bench on N GPUs (gpurun --gpus N)
mkdir -p gpurun_out
N=${1:-2}; TAG=${2:-dp}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/dp_check.py > gpurun_out/${TAG}_dp_check_n$N.log 2>&1; tail -3 gpurun_out/${TAG}_dp_check_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 200 --warmup 10 --cpu-steps 2 --cpu-queries 4 --no-c5 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err; tail -c 400 gpurun_out/${TAG}_bench_n$N.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/${TAG}_bench_n$N.json') if l.startswith('{')][-1])
print('N', d['n_gpus'], 'value', d['value'], 'ms', d['ms_per_step'], 'plain', d.get('value_plain_calls'), 'e2e', d['e2e']['value'], d['exchange'], d['step_api'])
print('graph', d.get('graph_phase_ms_per_step')); print(d['roofline']['phase_ms_per_step'], 'eval', d['eval']['value'], d['eval']['e2e']['value'])
PY
