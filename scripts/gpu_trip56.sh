#!/bin/bash
# N-GPU bench line exactly as the driver launches it
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r1g_bench_${N}gpu.json 2> gpurun_out/r1g_bench_${N}gpu.err
tail -2 gpurun_out/r1g_bench_${N}gpu.err | cut -c1-300
python - <<PY
import json
d = json.loads(open('gpurun_out/r1g_bench_${N}gpu.json').read().strip().splitlines()[-1])
print('MC ${N}gpu', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['n_gpus'])
t = d['train']; print('train ${N}gpu', t['value'], t['ms_per_step'], 'e2e', t['e2e']['value'])
PY
