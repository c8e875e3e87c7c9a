# model API under torchrun (replicas bit-identical, sharded evaluate, w_private case) + kernel-level dp_check
mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 scripts/api_dp_check.py > gpurun_out/api_dp_check_n$N.log 2>&1; tail -6 gpurun_out/api_dp_check_n$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/dp_check.py > gpurun_out/dp_check_n$N.log 2>&1; tail -3 gpurun_out/dp_check_n$N.log
