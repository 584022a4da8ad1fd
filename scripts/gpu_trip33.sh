#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_train_kernels_gpu.py tests/test_training_gpu.py -m gpu -q --timeout 600 -x 2>&1 | tail -5
timeout 600 python scripts/train_report.py --time 16 512 2>&1 | head -3
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
   --log-file gpurun_out/trip33_train_launches.csv python scripts/train_prof_step.py > gpurun_out/trip33_ncu.log 2>&1
python scripts/summarize_launches.py gpurun_out/trip33_train_launches.csv 2>/dev/null | head -24
for i in 1 2; do timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload mc 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('MC', d['value'], d['ms_per_step'], d['clocks'])"; done
