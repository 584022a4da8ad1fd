#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decoder_concat -s 1 -c 1 \
    -o gpurun_out/r1b_concat2 -f python scripts/prof_concat.py 80 10 128 256 128 64 32 > gpurun_out/trip17_ncu.log 2>&1
tail -2 gpurun_out/trip17_ncu.log
