mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -8 > gpurun_out/nrw_pytest.log; tail -8 gpurun_out/nrw_pytest.log
timeout 300 python scripts/c1_phases.py 2>&1 | tail -3
