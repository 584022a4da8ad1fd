#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_gpu.py tests/test_model_gpu.py -m gpu -q --timeout 300 2>&1 | tail -8 > gpurun_out/trip13_tests.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_b8_r1b.csv > gpurun_out/trip13_bench.log 2>&1
tail -4 gpurun_out/trip13_tests.log; tail -1 gpurun_out/trip13_bench.log | cut -c1-400; cat gpurun_out/layers_b8_r1b.csv
