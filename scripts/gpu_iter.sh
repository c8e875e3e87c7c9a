# usage: bash scripts/gpu_iter.sh <tag>  — GPU tests + short bench without the C5 leg (iteration loop)
mkdir -p gpurun_out
TAG=${1:-it}
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -25 > gpurun_out/${TAG}_pytest.log; tail -4 gpurun_out/${TAG}_pytest.log
python bench.py --steps 200 --warmup 10 --cpu-steps 3 --cpu-queries 16 --no-c5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; tail -c 600 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_bench.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'warm', d['value_steady_state_l2_warm'])
print(d['roofline']['phase_ms_per_step']); print('graph', d.get('graph_phase_ms_per_step'))
print('eval', d['eval']['value'], d['eval']['e2e']['value'], d['eval']['cpu_baseline'].get('ranks_outside_tie_band'))
PY
