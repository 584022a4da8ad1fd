#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests/test_model_gpu.py tests/test_memops_gpu.py -m gpu -q --timeout 600 -x 2>&1 | tail -3
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload mc 2>/dev/null | tail -1 | cut -c1-220
} > gpurun_out/trip52.log 2>&1
cat gpurun_out/trip52.log
