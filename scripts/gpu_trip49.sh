#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
echo "--- pair on"; timeout 300 python scripts/check_batch_invariance.py 2>&1 | tail -4
echo "--- pair off"; VU_CONV_PAIR=0 timeout 300 python scripts/check_batch_invariance.py 2>&1 | tail -4
} > gpurun_out/trip49.log 2>&1
cat gpurun_out/trip49.log
