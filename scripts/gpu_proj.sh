# projected-table ranking: parity tests, probe per model + launch lists, then the C1..C4 table
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rank.py tests/test_gpu_golden.py -m gpu -q -x 2>&1 | tail -30 > gpurun_out/proj_pytest.log
tail -15 gpurun_out/proj_pytest.log
for m in TransR TransD SE TransH; do timeout 300 python scripts/proj_probe.py $m 2>&1 | tail -1; done | tee gpurun_out/proj_probe.jsonl
for m in TransR TransD; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'proj|rank_lp|qrel' --csv --log-file gpurun_out/proj_launches_$m.csv python scripts/proj_probe.py $m > /dev/null 2>&1
done
timeout 600 python scripts/bench_configs.py > gpurun_out/proj_configs.jsonl 2> gpurun_out/proj_configs.err
tail -2 gpurun_out/proj_configs.jsonl | cut -c1-400; tail -5 gpurun_out/proj_configs.err
