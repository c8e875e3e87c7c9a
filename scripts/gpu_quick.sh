mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_rank.py tests/test_gpu_golden.py -m gpu -q -x 2>&1 | tail -15 > gpurun_out/pytest_quick.log
tail -3 gpurun_out/pytest_quick.log
python bench.py --steps 100 --warmup 5 --cpu-steps 3 --cpu-queries 16 > gpurun_out/quick_bench.json 2> gpurun_out/quick_bench.err; tail -c 400 gpurun_out/quick_bench.err
