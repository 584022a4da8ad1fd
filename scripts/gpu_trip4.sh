#!/bin/bash
# Full GPU check: all -m gpu tests, smoke, bench (plain), then an ncu launch list of a short bench.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 2>&1 | tail -60 > gpurun_out/trip4_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/trip4_smoke.log 2>&1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/trip4_bench.log 2>&1
echo "bench rc=$?" >> gpurun_out/trip4_bench.log
timeout 900 python bench.py --steps 10 --warmup 3 --dtype f16 --no-cpu-baseline > gpurun_out/trip4_bench_f16.log 2>&1
tail -30 gpurun_out/trip4_tests.log; tail -5 gpurun_out/trip4_smoke.log; tail -5 gpurun_out/trip4_bench.log
