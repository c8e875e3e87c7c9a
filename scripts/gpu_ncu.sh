# usage: bash scripts/gpu_ncu.sh <tag> <kernel-regex> [count] — ncu --set full capture of matching kernels in a short bench run
mkdir -p gpurun_out
TAG=${1:-run}; RGX=${2:-k_step}; CNT=${3:-3}
python bench.py --steps 6 --warmup 3 --cpu-steps 1 --cpu-queries 2 --no-c5 > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"$RGX" -c $CNT -o gpurun_out/${TAG}_prof python bench.py --steps 6 --warmup 3 --cpu-steps 1 --cpu-queries 2 --no-c5 > gpurun_out/${TAG}_ncu.log 2>&1
tail -3 gpurun_out/${TAG}_ncu.log
