#!/bin/bash
# round-1i: ncu launch lists of the FINAL kernels of the round (training + MC), same commands as the bench, no bench number taken here
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1i_train_launches.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload train > gpurun_out/r1i_ncu_train.log 2>&1
python scripts/summarize_launches.py gpurun_out/r1i_train_launches.csv > gpurun_out/r1i_train_launch_summary.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1i_mc_launches.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mc > gpurun_out/r1i_ncu_mc.log 2>&1
python scripts/mc_launch_summary.py gpurun_out/r1i_mc_launches.csv > gpurun_out/r1i_mc_launch_summary.txt 2>&1
head -30 gpurun_out/r1i_train_launch_summary.txt
head -14 gpurun_out/r1i_mc_launch_summary.txt
