#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_memops_gpu.py -m gpu -q --timeout 300 -x -k "staged or uncertainty" 2>&1 | tail -3
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2> gpurun_out/trip38.err | tail -1 > gpurun_out/trip38_bench.json
tail -3 gpurun_out/trip38.err
python -c "
import json
d = json.loads(open('gpurun_out/trip38_bench.json').read())
print('MC', d['value'], d['ms_per_step'], 'e2e', d['e2e'])
t = d['train']; print('train', t['value'], t['ms_per_step'], 'e2e', t['e2e'], t['gpu_launches'])
"
