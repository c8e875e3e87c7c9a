# projected-table ranking: parity tests, then the C1..C4 table
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rank.py tests/test_gpu_golden.py -m gpu -q -x 2>&1 | tail -30 > gpurun_out/proj_pytest.log
tail -15 gpurun_out/proj_pytest.log
timeout 600 python scripts/bench_configs.py > gpurun_out/proj_configs.jsonl 2> gpurun_out/proj_configs.err
cat gpurun_out/proj_configs.jsonl | cut -c1-600; tail -5 gpurun_out/proj_configs.err
