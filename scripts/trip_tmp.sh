#!/bin/bash
# scratch trip: filter-row weight-gradient threshold under the cost-model K split
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
O=gpurun_out/$1
for env in "VU_X=0" "VU_WGRAD_ROW3_MIN=512" "VU_WGRAD_ROW3_MIN=128" "VU_WGRAD_ROW3=1" "VU_WGRAD_BN256=1"; do
  echo "== $env"
  env $env python scripts/train_breakdown.py 2>&1 | tail -1
  env $env python scripts/train_breakdown.py 2>&1 | tail -1
done > ${O}_row3min.log 2>&1
cat ${O}_row3min.log
