# usage: bash scripts/gpu_bench_ncu.sh <tag> [kernel-regex]
mkdir -p gpurun_out
TAG=${1:-run}; RGX=${2:-k_rank_dot_tc}
python bench.py --steps 200 --warmup 10 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
tail -c 600 gpurun_out/${TAG}_bench.err
python bench.py --steps 6 --warmup 3 --cpu-steps 1 --cpu-queries 2 > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"$RGX" -c 4 -o gpurun_out/${TAG}_prof python bench.py --steps 6 --warmup 3 --cpu-steps 1 --cpu-queries 2 > gpurun_out/${TAG}_ncu.log 2>&1
tail -3 gpurun_out/${TAG}_ncu.log
