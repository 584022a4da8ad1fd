#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/trip37_mc_launches.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mc > gpurun_out/trip37_ncu.log 2>&1
python scripts/mc_launch_summary.py gpurun_out/trip37_mc_launches.csv --list | head -100
