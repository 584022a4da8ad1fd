#!/bin/bash
# 2-GPU check of the CTA-pair kernels next to NCCL: DP training parity, sharded MC parity, bench --gpus 2
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/dist_train_parity.py 16 512 2>&1 | grep -v Warning | tail -3 | tee gpurun_out/r1e_dp2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 scripts/dist_parity.py 2>&1 | grep -v Warning | tail -4 | tee -a gpurun_out/r1e_dp2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r1e_bench_2gpu.json 2> gpurun_out/r1e_bench2.err
tail -2 gpurun_out/r1e_bench2.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r1e_bench_2gpu.json').read().strip().splitlines()[-1])
print('MC 2gpu', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])
t = d['train']; print('train 2gpu', t['value'], t['ms_per_step'], 'e2e', t['e2e']['value'])
PY
