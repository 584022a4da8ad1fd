#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
timeout 300 python scripts/prof_bilinear_bwd.py 2>&1 | tail -8
timeout 600 python -m pytest tests/test_train_kernels_gpu.py tests/test_training_gpu.py -m gpu -q --timeout 300 -x 2>&1 | tail -2
timeout 600 python scripts/train_report.py --time 16 512 2>&1 | sed -n 2,3p
} > gpurun_out/trip46.log 2>&1
cat gpurun_out/trip46.log
