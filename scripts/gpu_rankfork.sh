#!/bin/bash
# A/B of the two-stream front of the tensor-core ranking path (KGE_RANK_SERIAL=1: one stream)
timeout 600 python -m pytest tests/test_gpu_rank.py tests/test_gpu_filter.py tests/test_gpu_fullsize.py -x -q -m gpu 2>&1 | tail -2
for v in 1 0 1 0; do
  KGE_RANK_SERIAL=$v python bench.py --steps 50 --warmup 5 --cpu-steps 1 --cpu-queries 8 --no-c5 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); e = d['eval']
print('SERIAL=$v', 'eval M q/s', round(e['value']/1e6, 2), 'e2e', round(e['e2e']['value']/1e6, 2), e['e2e']['device_ms_per_call'], 'outside band', e['cpu_baseline']['ranks_outside_tie_band'], 'identical', e['cpu_baseline']['ranks_identical'])
"
done
