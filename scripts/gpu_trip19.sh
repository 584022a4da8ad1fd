#!/bin/bash
# 2-GPU trip: NCCL parity of the sharded paths, then the bench at N=1 and N=2
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    scripts/dist_parity.py > gpurun_out/dist_parity_n$N.log 2>&1
grep -E "PASS|FAIL|Error|error" gpurun_out/dist_parity_n$N.log | head
timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.log 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.log 2>&1
tail -1 gpurun_out/bench_n1.log | cut -c1-160; tail -1 gpurun_out/bench_n$N.log | cut -c1-160
