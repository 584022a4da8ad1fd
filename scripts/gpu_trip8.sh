#!/bin/bash
# halo conv: parity in both descriptor modes, then standalone timing of the layers it takes over
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for m in 0 1; do
  echo "=== VU_HALO_BASEOFF=$m" >> gpurun_out/trip8_tests.log
  VU_HALO_BASEOFF=$m timeout 600 python -m pytest tests/test_conv_gpu.py -m gpu -q --timeout 120 -k "halo or persistent" 2>&1 | tail -25 >> gpurun_out/trip8_tests.log
done
for cfg in "16 256 256 224 64" "16 256 256 64 64" "16 128 128 352 128" "16 128 128 128 128" "8 128 128 64 64"; do
  for nh in 1 0; do
    echo "VU_NO_HALO=$nh" >> gpurun_out/trip8_perf.log
    VU_NO_HALO=$nh timeout 120 python scripts/prof_conv.py $cfg 3 4 >> gpurun_out/trip8_perf.log 2>&1
  done
done
cat gpurun_out/trip8_tests.log | grep -E "===|passed|failed|Error|error" | head; cat gpurun_out/trip8_perf.log
