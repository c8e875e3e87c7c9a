mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_step_rot' -c 1 --launch-skip 3 -o gpurun_out/rot_prof python scripts/phases_cfg.py RotatE > gpurun_out/rot_ncu.log 2>&1
tail -3 gpurun_out/rot_ncu.log
