#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests/test_conv_gpu.py -m gpu -q --timeout 120 -x 2>&1 | tail -3
for shape in "16 32 32 256 256 3" "16 64 64 128 128 3" "16 16 16 512 512 3" "80 32 32 800 512 3" "80 64 64 672 256 3" "16 128 128 128 352 3"; do
  echo "== $shape"
  timeout 120 python scripts/prof_conv.py $shape 6 bf16 2>&1 | tail -1
  VU_CONV_PAIR=0 timeout 120 python scripts/prof_conv.py $shape 6 bf16 2>&1 | tail -1
done
echo "--- pair on"
timeout 600 python scripts/train_report.py --time 16 512 2>&1 | sed -n 2,3p
echo "--- MC pair on"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --workload mc 2>/dev/null | tail -1 | cut -c1-200
} > gpurun_out/trip44.log 2>&1
cat gpurun_out/trip44.log
