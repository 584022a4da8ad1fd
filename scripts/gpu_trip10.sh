#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out; rm -f gpurun_out/trip10_perf.log
timeout 600 python -m pytest tests/test_conv_gpu.py -m gpu -q --timeout 120 -k "halo or persistent" 2>&1 | tail -5 > gpurun_out/trip10_tests.log
for cfg in "16 256 256 224 64" "16 256 256 64 64" "16 128 128 352 128" "16 128 128 128 128"; do
  for ht in 2 4; do
    echo "VU_HALO_HT=$ht" >> gpurun_out/trip10_perf.log
    VU_HALO_HT=$ht timeout 120 python scripts/prof_conv.py $cfg 3 4 >> gpurun_out/trip10_perf.log 2>&1
  done
done
cat gpurun_out/trip10_tests.log gpurun_out/trip10_perf.log
