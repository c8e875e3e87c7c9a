# matrix-model step kernels: parity tests, phase split
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_step.py tests/test_gpu_golden.py tests/test_gpu_fullsize.py tests/test_gpu_api.py tests/test_gpu_graph.py tests/test_gpu_dp.py tests/test_gpu_optim.py -m gpu -q -x 2>&1 | tail -30 > gpurun_out/mat_pytest.log
tail -6 gpurun_out/mat_pytest.log
for m in TransR SE; do timeout 200 python scripts/phases_cfg.py $m 2>&1 | tail -1; KGE_MAT_NO_RECX=1 timeout 200 python scripts/phases_cfg.py $m 2>&1 | tail -1; done | tee gpurun_out/mat_phases.jsonl
