#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
for i in 1 2 3; do
  echo "fused:    $(timeout 600 python scripts/train_report.py --time 16 512 2>&1 | grep 'step')"
  echo "separate: $(VU_FUSE_BN_STATS=0 timeout 600 python scripts/train_report.py --time 16 512 2>&1 | grep 'step')"
done
for cfg in "8 64" "4 128" "8 128"; do
  echo "--- fused $cfg"; timeout 900 python scripts/train_report.py $cfg 2>&1 | grep "=="
  echo "--- separate $cfg"; VU_FUSE_BN_STATS=0 timeout 900 python scripts/train_report.py $cfg 2>&1 | grep "=="
done
} > gpurun_out/trip41.log 2>&1
cat gpurun_out/trip41.log
