#!/bin/bash
# Full GPU check: all -m gpu tests, smoke, bench (plain), then an ncu launch list of a short bench.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 2>&1 | tail -60 > gpurun_out/trip3_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/trip3_smoke.log 2>&1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/trip3_bench.log 2>&1
echo "bench rc=$?" >> gpurun_out/trip3_bench.log
timeout 600 python bench.py --steps 1 --warmup 3 --batch 2 --no-cpu-baseline > gpurun_out/trip3_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 500 --csv \
    --log-file gpurun_out/trip3_launches.csv python bench.py --steps 1 --warmup 3 --batch 2 --no-cpu-baseline > gpurun_out/trip3_ncu.log 2>&1
tail -30 gpurun_out/trip3_tests.log; tail -5 gpurun_out/trip3_smoke.log; tail -5 gpurun_out/trip3_bench.log
