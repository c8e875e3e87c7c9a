mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_step_proj|k_update_mat' -c 4 -o gpurun_out/mat_prof python scripts/phases_cfg.py TransR > gpurun_out/mat_ncu.log 2>&1
tail -3 gpurun_out/mat_ncu.log
