mkdir -p gpurun_out
python scripts/bench_c5.py --profile-only --B 8192 > gpurun_out/c5_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -c 300 --csv --log-file gpurun_out/c5_launches.csv python scripts/bench_c5.py --profile-only --B 8192 > gpurun_out/c5_ncu.log 2>&1
tail -2 gpurun_out/c5_ncu.log
