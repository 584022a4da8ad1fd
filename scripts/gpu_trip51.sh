#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decoder_concat -s 2 -c 1 -o gpurun_out/r1e_concat_b3 \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload mc > gpurun_out/trip51_ncu.log 2>&1
tail -2 gpurun_out/trip51_ncu.log
