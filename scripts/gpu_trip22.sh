#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python scripts/train_report.py --all 8 64 > gpurun_out/trip22_report.log 2>&1
timeout 900 python scripts/train_report.py --all 8 128 > gpurun_out/trip22_report128.log 2>&1
grep "==" gpurun_out/trip22_report.log gpurun_out/trip22_report128.log | cut -c1-400
