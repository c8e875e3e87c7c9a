mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dp.py tests/test_gpu_step.py -m gpu -q -x 2>&1 | tail -8
