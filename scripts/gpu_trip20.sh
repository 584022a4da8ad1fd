#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_train_kernels_gpu.py -m gpu -q --timeout 300 -x 2>&1 | tail -30 > gpurun_out/trip20_tests.log
cat gpurun_out/trip20_tests.log
