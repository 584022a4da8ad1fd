#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests/test_train_kernels_gpu.py tests/test_training_gpu.py tests/test_memops_gpu.py -m gpu -q --timeout 300 -x 2>&1 | tail -3
timeout 600 python scripts/train_report.py --time 16 512 2>&1 | sed -n 2,3p
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/trip45_train_launches.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload train > gpurun_out/trip45_ncu.log 2>&1
python scripts/summarize_launches.py gpurun_out/trip45_train_launches.csv | head -40
} > gpurun_out/trip45.log 2>&1
cat gpurun_out/trip45.log
