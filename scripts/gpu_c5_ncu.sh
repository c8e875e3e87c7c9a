# C5 (true HBM case): plain run, then one ncu --set full capture of the row-lazy step's kernels at B = 65536
mkdir -p gpurun_out
python scripts/bench_c5.py --profile-only --B 65536 > gpurun_out/c5_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_lazy_pre|k_lazy_update|k_step' --launch-skip 9 -c 6 -o gpurun_out/r01_c5_prof \
  python scripts/bench_c5.py --profile-only --B 65536 > gpurun_out/c5_ncu.log 2>&1
tail -2 gpurun_out/c5_ncu.log; cat gpurun_out/c5_plain.log | cut -c1-300
