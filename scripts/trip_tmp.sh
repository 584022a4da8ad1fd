#!/bin/bash
# scratch trip: K-split policy of the weight gradients, in-situ kernel sums and step time
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
O=gpurun_out/$1
python -m pytest tests/test_train_kernels_gpu.py tests/test_guards_gpu.py -x -q -m gpu -k "wgrad" 2>&1 | tail -2
for env in "VU_WGRAD_HALFWAVES=3" "VU_WGRAD_OVERHEAD=3" "VU_WGRAD_OVERHEAD=6" "VU_WGRAD_OVERHEAD=12" "VU_WGRAD_OVERHEAD=24"; do
  echo "== $env"
  env $env python scripts/train_kineto.py 2>&1 | grep "span\|wgrad"
  env $env python scripts/train_breakdown.py 2>&1 | tail -1
done > ${O}_ksplit.log 2>&1
cat ${O}_ksplit.log
