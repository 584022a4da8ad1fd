#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out; rm -f gpurun_out/parity_report.txt
timeout 900 python -m pytest tests -m gpu -q --timeout 300 2>&1 | tail -15 > gpurun_out/trip15_tests.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/trip15_bench.log 2>&1
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --batch 16 > gpurun_out/trip15_bench16.log 2>&1
tail -12 gpurun_out/trip15_tests.log; tail -1 gpurun_out/trip15_bench.log | cut -c1-200;  tail -1 gpurun_out/trip15_bench16.log | cut -c1-200
