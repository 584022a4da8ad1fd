#!/bin/bash
# scratch trip: flat padded tap-reuse kernel vs the CTA-pair implicit GEMM, layer shapes of the training step (N=16) and MC encoder
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
O=gpurun_out/$1
for shape in "16 64 64 128 128 3" "16 32 32 256 256 3" "16 16 16 512 512 3" "16 32 32 800 512 3" "16 32 32 512 512 3" "16 64 64 672 256 3" "16 64 64 256 256 3" "16 64 64 320 256 3"; do
  for env in "VU_FLAT=0" "VU_FLAT=2 VU_FLAT_MT=2" "VU_FLAT=2 VU_FLAT_MT=1"; do
    echo -n "$env  "; env $env python scripts/prof_conv.py $shape 6 f16 | tail -1
    echo -n "$env stats "; env $env python scripts/prof_conv.py $shape 6 f16 stats | tail -1
  done
done > ${O}_flat_vs_pair.log 2>&1
cat ${O}_flat_vs_pair.log
