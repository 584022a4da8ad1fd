#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
echo "--- fused" > gpurun_out/trip39.log
timeout 900 python -m pytest tests/test_training_gpu.py -m gpu -q --timeout 600 2>&1 | tail -60 >> gpurun_out/trip39.log
echo "--- separate" >> gpurun_out/trip39.log
VU_FUSE_BN_STATS=0 timeout 900 python -m pytest tests/test_training_gpu.py -m gpu -q --timeout 600 2>&1 | tail -60 >> gpurun_out/trip39.log
echo "--- fused stats" >> gpurun_out/trip39.log
timeout 600 python scripts/train_report.py --time 16 512 >> gpurun_out/trip39.log 2>&1 
echo "--- separate stats" >> gpurun_out/trip39.log
VU_FUSE_BN_STATS=0 timeout 600 python scripts/train_report.py --time 16 512 >> gpurun_out/trip39.log 2>&1
tail -3 gpurun_out/trip39.log
