#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_train_kernels_gpu.py tests/test_training_gpu.py tests/test_conv_gpu.py -m gpu -q --timeout 600 -x 2>&1 | tail -15 > gpurun_out/trip24_tests.log
timeout 600 python scripts/train_report.py --time 16 512 > gpurun_out/trip24_time.log 2>&1
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
   --log-file gpurun_out/trip24_train_launches.csv python scripts/train_prof_step.py > gpurun_out/trip24_ncu.log 2>&1
tail -8 gpurun_out/trip24_tests.log; head -4 gpurun_out/trip24_time.log
python scripts/summarize_launches.py gpurun_out/trip24_train_launches.csv 2>/dev/null | head -36
