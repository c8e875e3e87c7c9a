mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/t1_pytest.log; tail -3 gpurun_out/t1_pytest.log
timeout 600 python scripts/bench_configs.py > gpurun_out/t1_configs.jsonl 2> gpurun_out/t1_configs.err; tail -c 300 gpurun_out/t1_configs.err; cut -c1-260 gpurun_out/t1_configs.jsonl
