#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_gpu.py tests/test_model_gpu.py tests/test_memops_gpu.py -m gpu -q --timeout 600 -x 2>&1 | tail -8
for i in 1 2; do timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload mc 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('MC', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['roofline']['frac'], d['clocks'])"; done
grep "mc_predict f16\|forward f16" gpurun_out/parity_report.txt | tail -5
