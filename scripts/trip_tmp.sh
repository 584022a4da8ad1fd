#!/bin/bash
# scratch trip: DP training with the NCCL CTA limit vs the default communicator (2 GPUs)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for c in 2 0 1; do
  VU_DP_NCCL_CTAS=$c python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $((29800+c)) \
    bench.py --gpus $2 --workload train --no-cpu-baseline > gpurun_out/$1_train_ctas$c.json 2> gpurun_out/$1_train_ctas$c.err
  echo "ctas=$c"; grep '^{' gpurun_out/$1_train_ctas$c.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])"; tail -2 gpurun_out/$1_train_ctas$c.err | cut -c1-200
done
