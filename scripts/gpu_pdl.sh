#!/bin/bash
# A/B of programmatic dependent launch of the row update behind the step kernel (KGE_PDL=0 / 1)
set -x
timeout 600 python -m pytest tests/test_gpu_step.py tests/test_gpu_graph.py tests/test_gpu_golden.py -x -q -m gpu 2>&1 | tail -3
for v in 0 1 0 1; do
  KGE_PDL=$v python bench.py --steps 200 --warmup 10 --cpu-steps 1 --cpu-queries 4 --no-c5 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('PDL=$v', 'us/step', round(d['ms_per_step']*1e3, 2), 'warm', round(d['value_steady_state_l2_warm']/1e6, 1), 'plain', round(d['value_plain_calls']/1e6, 1), 'e2e', round(d['e2e']['value']/1e6, 1))
"
done
