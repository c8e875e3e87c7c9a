mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 100 --warmup 5 --cpu-steps 2 --cpu-queries 4 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; tail -c 300 gpurun_out/bench_n$N.err; cut -c1-900 gpurun_out/bench_n$N.json
