mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_step.py tests/test_gpu_fullsize.py tests/test_gpu_lazy.py tests/test_gpu_dp.py tests/test_gpu_graph.py -m gpu -q -x 2>&1 | tail -6
timeout 300 python scripts/c1_phases.py 2>&1 | tail -3
