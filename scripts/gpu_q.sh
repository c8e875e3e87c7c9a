mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/q_pytest.log; grep -n "Error\|assert \|FAILED" gpurun_out/q_pytest.log | head -20; tail -3 gpurun_out/q_pytest.log
