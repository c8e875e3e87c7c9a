mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/t6_pytest.log; grep -n "Error\|assert \|FAILED" gpurun_out/t6_pytest.log | head -10; tail -2 gpurun_out/t6_pytest.log
for m in TransR SE RESCAL; do python scripts/phases_cfg.py $m 2>&1 | tail -1; done
