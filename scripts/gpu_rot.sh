mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_step.py tests/test_gpu_golden.py tests/test_gpu_fullsize.py tests/test_gpu_lazy.py -m gpu -q -x 2>&1 | tail -30 > gpurun_out/rot_pytest.log
tail -12 gpurun_out/rot_pytest.log
timeout 200 python scripts/phases_cfg.py RotatE 2>&1 | tail -1; KGE_STEP_ROT_OLD=1 timeout 200 python scripts/phases_cfg.py RotatE 2>&1 | tail -1
