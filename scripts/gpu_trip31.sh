#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_train_kernels_gpu.py -m gpu -q --timeout 300 -x 2>&1 | tail -12 > gpurun_out/trip31_kernels.log
tail -6 gpurun_out/trip31_kernels.log
rm -f gpurun_out/trip31_wgrad.log
for shp in "16 32 32 256 256 3" "16 128 128 64 64 3" "16 256 256 224 64 3" "16 64 64 128 128 3" "16 16 16 512 512 3" "16 256 256 64 64 3" "16 128 128 352 128 3"; do
  timeout 120 python scripts/prof_wgrad.py $shp 6 >> gpurun_out/trip31_wgrad.log 2>&1
  VU_WGRAD_PER_TAP=1 timeout 120 python scripts/prof_wgrad.py $shp 6 2>&1 | sed 's/^/   per-tap: /' >> gpurun_out/trip31_wgrad.log
done
cat gpurun_out/trip31_wgrad.log
timeout 900 python -m pytest tests/test_training_gpu.py tests/test_conv_gpu.py -m gpu -q --timeout 600 -x 2>&1 | tail -6
timeout 600 python scripts/train_report.py --time 16 512 2>&1 | head -14
