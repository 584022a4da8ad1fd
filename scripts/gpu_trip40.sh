#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
for cfg in "2 64" "8 64" "4 256"; do
  echo "--- fused $cfg"; timeout 600 python scripts/check_run_grads.py $cfg 2>&1 | grep -A1 "run 0 vs"
  echo "--- separate $cfg"; VU_FUSE_BN_STATS=0 timeout 600 python scripts/check_run_grads.py $cfg 2>&1 | grep -A1 "run 0 vs"
done
timeout 900 python -m pytest tests/test_training_gpu.py tests/test_train_kernels_gpu.py -m gpu -q --timeout 600 2>&1 | tail -5
timeout 900 python -m pytest tests/test_training_gpu.py -m gpu -q --timeout 600 2>&1 | tail -2
timeout 900 python -m pytest tests/test_training_gpu.py -m gpu -q --timeout 600 2>&1 | tail -2
timeout 600 python scripts/train_report.py --time 16 512 2>&1 | grep "step"
} > gpurun_out/trip40.log 2>&1
cat gpurun_out/trip40.log
