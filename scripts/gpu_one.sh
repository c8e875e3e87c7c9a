# usage: bash scripts/gpu_one.sh <tag> <pytest args...> — selected GPU tests
mkdir -p gpurun_out
TAG=$1; shift
timeout 900 python -m pytest "$@" -m gpu -q -x 2>&1 | tail -40 > gpurun_out/${TAG}_pytest.log; tail -30 gpurun_out/${TAG}_pytest.log
