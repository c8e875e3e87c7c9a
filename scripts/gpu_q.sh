mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -60 > gpurun_out/q_pytest.log; grep -n "Error\|error\|assert\|FAILED" gpurun_out/q_pytest.log | head -20; tail -3 gpurun_out/q_pytest.log
