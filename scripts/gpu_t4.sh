mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/t4_pytest.log; grep -n "Error\|assert \|FAILED" gpurun_out/t4_pytest.log | head -10; tail -2 gpurun_out/t4_pytest.log
timeout 600 python scripts/bench_configs.py > gpurun_out/t4_configs.jsonl 2> gpurun_out/t4_configs.err; tail -c 300 gpurun_out/t4_configs.err; cut -c1-190 gpurun_out/t4_configs.jsonl
timeout 600 python scripts/bench_c5.py --profile-only --B 65536 2>&1 | tail -1 | cut -c1-250
timeout 600 python scripts/bench_c5.py --profile-only --B 8192 2>&1 | tail -1 | cut -c1-250
python bench.py --steps 200 --warmup 10 --cpu-steps 2 --cpu-queries 8 > gpurun_out/t4_bench.json 2> gpurun_out/t4_bench.err; tail -c 300 gpurun_out/t4_bench.err
python -c "
import json; d=json.load(open('gpurun_out/t4_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['phase_ms_per_step'], d['roofline']['frac'])"
