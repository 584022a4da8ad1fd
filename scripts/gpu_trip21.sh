#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_train_kernels_gpu.py -m gpu -q --timeout 300 -x 2>&1 | tail -15 > gpurun_out/trip21_kernels.log
timeout 600 python scripts/train_report.py > gpurun_out/trip21_report.log 2>&1
timeout 600 python scripts/train_report.py --time 16 512 > gpurun_out/trip21_time.log 2>&1
tail -5 gpurun_out/trip21_kernels.log; tail -60 gpurun_out/trip21_report.log; tail -50 gpurun_out/trip21_time.log
