mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/t2_pytest.log; tail -2 gpurun_out/t2_pytest.log
timeout 600 python scripts/bench_configs.py 2> gpurun_out/t2_configs.err | head -2 > gpurun_out/t2_configs.jsonl; cut -c1-200 gpurun_out/t2_configs.jsonl
timeout 900 python scripts/bench_c5.py --dense-steps 3 --steps 60 > gpurun_out/c5_v3.jsonl 2> gpurun_out/c5_v3.err; tail -c 600 gpurun_out/c5_v3.err; grep -v dense gpurun_out/c5_v3.jsonl | cut -c1-420
