#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests/test_training_gpu.py tests/test_train_kernels_gpu.py -m gpu -q --timeout 600 -x 2>&1 | tail -2
for i in 1 2; do timeout 600 python scripts/train_report.py --time 16 512 2>&1 | sed -n 2,2p; done
} > gpurun_out/trip62.log 2>&1
cat gpurun_out/trip62.log
