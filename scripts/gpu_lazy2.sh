mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_shard.py tests/test_gpu_lazy.py tests/test_gpu_step.py tests/test_gpu_dp.py -m gpu -q -x 2>&1 | tail -8 > gpurun_out/lazy2_pytest.log; tail -3 gpurun_out/lazy2_pytest.log
timeout 900 python scripts/bench_c5.py --dense-steps 3 --steps 60 > gpurun_out/c5_v2.jsonl 2> gpurun_out/c5_v2.err; tail -c 600 gpurun_out/c5_v2.err; grep -v dense gpurun_out/c5_v2.jsonl | cut -c1-520
