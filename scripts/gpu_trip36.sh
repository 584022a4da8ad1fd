#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
for i in 1 2; do timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload mc 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('MC graph', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['gpu_launches'], d['roofline']['frac'])"; done
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload mc --no-graph 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('MC nograph', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['gpu_launches'])"
