#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
echo "--- all switches off"
VU_CONV_PAIR=0 VU_HALO_PAIR=0 VU_FUSE_BN_STATS=0 timeout 1500 python -m pytest tests -m gpu -q --timeout 900 2>&1 | tail -2
echo "--- run-to-run reproducibility of the training gradient (default switches)"
timeout 600 python scripts/check_run_grads.py 8 64 2>&1 | grep -A1 "run 0 vs"
timeout 600 python scripts/check_run_grads.py 16 256 2>&1 | grep -A1 "run 0 vs"
} > gpurun_out/trip59.log 2>&1
cat gpurun_out/trip59.log
