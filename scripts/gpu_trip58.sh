#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
for i in 1 2 3; do
for u in 1 2; do
echo -n "unroll $u: "; VU_CAT_UNROLL=$u timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload mc 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['roofline']['frac'])"
done
done
} > gpurun_out/trip58.log 2>&1
cat gpurun_out/trip58.log
