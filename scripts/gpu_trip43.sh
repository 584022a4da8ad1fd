#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests/test_conv_gpu.py -m gpu -q --timeout 120 -x 2>&1 | tail -8
echo "--- pair on"
timeout 600 python scripts/train_report.py --time 16 512 2>&1 | head -16
echo "--- pair off"
VU_CONV_PAIR=0 timeout 600 python scripts/train_report.py --time 16 512 2>&1 | head -16
echo "--- MC pair on"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --workload mc 2>/dev/null | tail -1 | cut -c1-200
echo "--- MC pair off"
VU_CONV_PAIR=0 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --workload mc 2>/dev/null | tail -1 | cut -c1-200
} > gpurun_out/trip43.log 2>&1
cat gpurun_out/trip43.log
