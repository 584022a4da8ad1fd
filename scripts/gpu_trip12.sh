#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv3x3_halo -s 2 -c 1 \
    -o gpurun_out/r1b_halo_64to64 -f python scripts/prof_conv.py 16 256 256 64 64 3 4 > gpurun_out/trip12_a.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv3x3_halo -s 2 -c 1 \
    -o gpurun_out/r1b_halo_224to64 -f python scripts/prof_conv.py 16 256 256 224 64 3 4 > gpurun_out/trip12_b.log 2>&1
tail -3 gpurun_out/trip12_a.log gpurun_out/trip12_b.log
