#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 -x 2>&1 | tail -8 > gpurun_out/trip28_tests.log
tail -4 gpurun_out/trip28_tests.log
for b in 8 16 32; do
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --batch $b 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('batch', d['config']['patches_per_step_per_gpu'], 'value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'conv frac', d['roofline']['frac'])"
done
timeout 600 python scripts/train_report.py --time 16 512 2>&1 | head -3
