set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r1_pytest.log
python bench.py --steps 200 --warmup 10 > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err
python bench.py --steps 6 --warmup 3 --cpu-steps 1 > gpurun_out/r1_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_step|k_update' -s 6 -c 6 -o gpurun_out/r1_prof python bench.py --steps 6 --warmup 3 --cpu-steps 1 > gpurun_out/r1_ncu.log 2>&1
ls -la gpurun_out
