#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 -x 2>&1 | tail -4 > gpurun_out/trip30_tests.log
tail -2 gpurun_out/trip30_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/dist_train_parity.py 4 128 2>&1 | grep -v Warning | tail -3 | tee gpurun_out/trip30_dp.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 scripts/dist_train_parity.py 16 512 2>&1 | grep -v Warning | tail -3 | tee -a gpurun_out/trip30_dp.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/trip30_bench2.json 2> gpurun_out/trip30_bench2.err
python -c "
import json
d = json.loads(open('gpurun_out/trip30_bench2.json').read().strip().splitlines()[-1])
print('MC', d['value'], 'e2e', d['e2e']['value'], 'train', d['train']['value'], 'ms', d['train']['ms_per_step'], 'e2e', d['train']['e2e']['value'])
"
tail -3 gpurun_out/trip30_bench2.err
