# usage: bash scripts/gpu_full.sh <tag>  — GPU tests, bench, reference arm, launch list, one ncu --set full capture
mkdir -p gpurun_out
TAG=${1:-run}
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/${TAG}_pytest.log; tail -3 gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -1 gpurun_out/${TAG}_smoke.log
python bench.py --steps 200 --warmup 10 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; tail -c 300 gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
python bench.py --steps 6 --warmup 3 --cpu-steps 1 --cpu-queries 2 > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${TAG}_launches.csv \
  python bench.py --steps 6 --warmup 3 --cpu-steps 1 --cpu-queries 2 > gpurun_out/${TAG}_ncu_list.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_step|k_update_fast|k_canon' -c 6 -o gpurun_out/${TAG}_prof \
  python bench.py --steps 6 --warmup 3 --cpu-steps 1 --cpu-queries 2 > gpurun_out/${TAG}_ncu.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_rank_dot' -c 4 -o gpurun_out/${TAG}_prof_rank \
  python bench.py --steps 6 --warmup 3 --cpu-steps 1 --cpu-queries 2 > gpurun_out/${TAG}_ncu_rank.log 2>&1
tail -2 gpurun_out/${TAG}_ncu.log
timeout 600 python scripts/bench_configs.py > gpurun_out/${TAG}_configs.jsonl 2> gpurun_out/${TAG}_configs.err; wc -l gpurun_out/${TAG}_configs.jsonl
