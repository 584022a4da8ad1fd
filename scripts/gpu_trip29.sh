#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 -x 2>&1 | tail -6 > gpurun_out/trip29_tests.log
tail -3 gpurun_out/trip29_tests.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/trip29_bench.json 2> gpurun_out/trip29_bench.err
tail -c 3000 gpurun_out/trip29_bench.json; tail -3 gpurun_out/trip29_bench.err
