#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/r1b_launches_b8.csv python bench.py --steps 1 --warmup 3 --batch 8 --no-cpu-baseline > gpurun_out/trip14_ncu.log 2>&1
tail -2 gpurun_out/trip14_ncu.log | cut -c1-300; wc -l gpurun_out/r1b_launches_b8.csv
