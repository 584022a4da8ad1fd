#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out; rm -f gpurun_out/trip9_perf.log
for cfg in "16 256 256 224 64" "16 256 256 64 64"; do
  for ht in 2 4; do
    echo "VU_HALO_HT=$ht" >> gpurun_out/trip9_perf.log
    VU_HALO_HT=$ht timeout 120 python scripts/prof_conv.py $cfg 3 4 >> gpurun_out/trip9_perf.log 2>&1
  done
done
cat gpurun_out/trip9_perf.log
