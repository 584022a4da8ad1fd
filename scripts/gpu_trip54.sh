#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
VU_HALO_PAIR=1 timeout 300 python -m pytest tests/test_conv_gpu.py -m gpu -q --timeout 120 -x 2>&1 | tail -4
for shape in "80 256 256 64 64 3" "80 256 256 192 64 3" "80 128 128 128 128 3" "16 128 128 64 64 3"; do
  echo "== $shape"
  VU_HALO_PAIR=1 timeout 120 python scripts/prof_conv.py $shape 5 f16 2>&1 | tail -1
  timeout 120 python scripts/prof_conv.py $shape 5 f16 2>&1 | tail -1
done
echo "--- MC halo pair on"
VU_HALO_PAIR=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload mc 2>/dev/null | tail -1 | cut -c1-200
echo "--- MC halo pair off"
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload mc 2>/dev/null | tail -1 | cut -c1-200
} > gpurun_out/trip54.log 2>&1
cat gpurun_out/trip54.log
