# usage: bash scripts/gpu_ab.sh VAR v1 v2 ... — bench (no C5, short CPU legs) once per value of env VAR
mkdir -p gpurun_out
VAR=$1; shift
for V in "$@"; do
env $VAR=$V python bench.py --steps 200 --warmup 10 --cpu-steps 2 --cpu-queries 4 --no-c5 > gpurun_out/ab_${VAR}_$V.json 2> gpurun_out/ab_${VAR}_$V.err; tail -c 300 gpurun_out/ab_${VAR}_$V.err
python -c "
import json; d=json.loads(open('gpurun_out/ab_${VAR}_$V.json').read().strip().splitlines()[-1]); print('$VAR=$V', 'value', round(d['value']/1e6,1), 'us', round(d['ms_per_step']*1e3,2), 'e2e', round(d['e2e']['value']/1e6,1), 'warm', round(d['value_steady_state_l2_warm']/1e6,1), d['roofline']['phase_ms_per_step'], d['graph_phase_ms_per_step'])"
done
