mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q "$@" 2>&1 | tail -60 > gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
