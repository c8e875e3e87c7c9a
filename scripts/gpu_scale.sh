# usage: bash scripts/gpu_scale.sh N   — bench.py on N GPUs the way the driver launches it
mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 200 --warmup 10 --cpu-steps 2 --cpu-queries 4 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; tail -c 300 gpurun_out/scale_n$N.err
python -c "
import json; d=json.loads([l for l in open('gpurun_out/scale_n$N.json') if l.startswith('{')][-1]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['config']['exchange'], d['eval']['value'], d['clocks'])"
