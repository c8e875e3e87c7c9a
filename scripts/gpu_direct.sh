#!/bin/bash
# step graph with the batch / loss crossing the host boundary inside the graph (KGE_GRAPH_DIRECT_IO=0: copies around it)
timeout 600 python -m pytest tests/test_gpu_graph.py tests/test_gpu_api.py tests/test_gpu_dp.py -x -q -m gpu 2>&1 | tail -3
python scripts/e2e_graph_probe.py 2>&1 | tail -5
for v in 0 1; do
  KGE_GRAPH_DIRECT_IO=$v python bench.py --steps 200 --warmup 10 --cpu-steps 1 --cpu-queries 4 --no-c5 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('DIRECT=$v', 'us/step', round(d['ms_per_step']*1e3, 2), 'warm', round(d['value_steady_state_l2_warm']/1e6, 1), 'e2e', round(d['e2e']['value']/1e6, 1), 'late', round(d['e2e_loss_one_step_late']['value']/1e6, 1), 'loss', d['final_loss'])
"
done
