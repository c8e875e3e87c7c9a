mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 scripts/shard_check.py --bench > gpurun_out/shard_check_n$N.log 2>&1; tail -12 gpurun_out/shard_check_n$N.log | cut -c1-900
