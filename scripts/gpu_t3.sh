mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/t3_pytest.log; grep -n "Error\|assert \|FAILED" gpurun_out/t3_pytest.log | head -10; tail -2 gpurun_out/t3_pytest.log
python bench.py --steps 200 --warmup 10 --cpu-steps 2 --cpu-queries 8 > gpurun_out/t3_bench.json 2> gpurun_out/t3_bench.err; tail -c 300 gpurun_out/t3_bench.err
python -c "
import json; d=json.load(open('gpurun_out/t3_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['phase_ms_per_step'], d['roofline']['frac'])"
