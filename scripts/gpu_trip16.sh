#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_memops_gpu.py -m gpu -q --timeout 120 -k "concat" 2>&1 | tail -5 > gpurun_out/trip16.log
python scripts/prof_concat.py 80 10 128 256 128 64 32 >> gpurun_out/trip16.log 2>&1
python scripts/prof_concat.py 80 10 64 128 256 64 64 >> gpurun_out/trip16.log 2>&1
python scripts/prof_concat.py 80 10 32 64 512 128 128 >> gpurun_out/trip16.log 2>&1
cat gpurun_out/trip16.log
