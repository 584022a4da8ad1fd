#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
for b in 16 24 32; do
  echo "--- MC batch $b"
  timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --workload mc --batch $b 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['roofline']['frac'])"
done
} > gpurun_out/trip57.log 2>&1
cat gpurun_out/trip57.log
