mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lazy.py -m gpu -q -x 2>&1 | tail -30 > gpurun_out/lazy_pytest.log; tail -5 gpurun_out/lazy_pytest.log
timeout 900 python scripts/bench_c5.py "$@" > gpurun_out/c5.jsonl 2> gpurun_out/c5.err; tail -c 1500 gpurun_out/c5.err; cat gpurun_out/c5.jsonl | cut -c1-600
