# A/B of the TMA row ring depth on the bench (no tests): bash scripts/gpu_iter2.sh <tag>
mkdir -p gpurun_out
TAG=${1:-it}
for RS in 32 8 0; do
KGE_RING_SLOTS=$RS python bench.py --steps 200 --warmup 10 --cpu-steps 2 --cpu-queries 4 --no-c5 > gpurun_out/${TAG}_rs${RS}.json 2> gpurun_out/${TAG}_rs${RS}.err; tail -c 300 gpurun_out/${TAG}_rs${RS}.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_rs${RS}.json').read().strip().splitlines()[-1])
print('RS=$RS value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'warm', d['value_steady_state_l2_warm'], d['roofline']['phase_ms_per_step'], d['final_loss'])
PY
done
