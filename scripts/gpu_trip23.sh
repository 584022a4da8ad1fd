#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 -x 2>&1 | tail -15 > gpurun_out/trip23_tests.log
timeout 600 python scripts/train_prof_step.py > gpurun_out/trip23_prof_plain.log 2>&1 && \
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
   --log-file gpurun_out/trip23_train_launches.csv python scripts/train_prof_step.py > gpurun_out/trip23_ncu.log 2>&1
tail -5 gpurun_out/trip23_tests.log; tail -3 gpurun_out/trip23_ncu.log
python scripts/summarize_launches.py gpurun_out/trip23_train_launches.csv | head -50
