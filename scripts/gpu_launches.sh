# usage: bash scripts/gpu_launches.sh <tag>   — plain run first, then the ncu launch list of the same command
mkdir -p gpurun_out
TAG=${1:-run}
python bench.py --steps 6 --warmup 3 --cpu-steps 1 --cpu-queries 2 > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${TAG}_launches.csv \
  python bench.py --steps 6 --warmup 3 --cpu-steps 1 --cpu-queries 2 > gpurun_out/${TAG}_ncu_list.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_list.log
