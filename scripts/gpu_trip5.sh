#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 2>&1 | tail -40 > gpurun_out/trip5_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/trip5_smoke.log 2>&1
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_b8.csv > gpurun_out/trip5_bench.log 2>&1
tail -15 gpurun_out/trip5_tests.log; tail -5 gpurun_out/trip5_smoke.log; tail -2 gpurun_out/trip5_bench.log
