#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_train_kernels_gpu.py tests/test_training_gpu.py -m gpu -q --timeout 600 -x 2>&1 | tail -15 > gpurun_out/trip26_tests.log
for shp in "16 32 32 256 256 3" "16 128 128 64 64 3" "16 256 256 224 64 3" "16 64 64 128 128 3" "16 16 16 512 512 3"; do
  python scripts/prof_wgrad.py $shp 6 >> gpurun_out/trip26_wgrad.log 2>&1
done
timeout 600 python scripts/train_report.py --time 16 512 > gpurun_out/trip26_time.log 2>&1
tail -8 gpurun_out/trip26_tests.log; cat gpurun_out/trip26_wgrad.log; head -24 gpurun_out/trip26_time.log
