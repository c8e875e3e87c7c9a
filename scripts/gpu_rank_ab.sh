# ranking GEMM: tests + eval numbers with 3 (default) and 2 MMAs per K step
mkdir -p gpurun_out
for P in 3 2; do
KGE_RANK_TC_PASSES=$P timeout 600 python -m pytest tests/test_gpu_rank.py tests/test_gpu_fullsize.py -m gpu -q -x -k "tensor_core or rank" 2>&1 | tail -3
KGE_RANK_TC_PASSES=$P python bench.py --steps 50 --warmup 5 --cpu-steps 2 --cpu-queries 64 --no-c5 > gpurun_out/rk${P}_bench.json 2>gpurun_out/rk${P}_bench.err
python -c "
import json; d=json.loads(open('gpurun_out/rk${P}_bench.json').read().strip().splitlines()[-1]); e=d['eval']; print('passes $P', 'q/s', e['value'], 'e2e', e['e2e']['value'], 'TF', e['roofline']['achieved'], 'gemm ms', e['roofline']['avg_launch_ms'], 'fix ms', e['roofline']['fixup_filter_ms'], e['cpu_baseline']['ranks_identical'], e['cpu_baseline']['ranks_outside_tie_band'])"
done
