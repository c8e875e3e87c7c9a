mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_shard.py tests/test_gpu_lazy.py -m gpu -q -x 2>&1 | tail -40 > gpurun_out/shard_pytest.log; tail -25 gpurun_out/shard_pytest.log
