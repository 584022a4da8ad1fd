#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_model_gpu.py -m gpu -q --timeout 300 -s 2>&1 | tail -80 > gpurun_out/trip2_model.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/trip2_smoke.log 2>&1
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/trip2_bench.log 2>&1
echo "bench rc=$?" >> gpurun_out/trip2_bench.log
tail -40 gpurun_out/trip2_model.log; cat gpurun_out/trip2_smoke.log | tail -5; tail -5 gpurun_out/trip2_bench.log
