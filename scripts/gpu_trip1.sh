#!/bin/bash
# First GPU trip: each conv case in its own process (a trap poisons the CUDA context), then memops.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python - <<'PY' > gpurun_out/conv_cases.txt
import sys; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import test_conv_gpu as t
print("\n".join(c[0] for c in t.CASES))
PY
: > gpurun_out/trip1_conv.log
for c in $(cat gpurun_out/conv_cases.txt); do
  echo "=== $c" >> gpurun_out/trip1_conv.log
  timeout 180 python -m pytest tests/test_conv_gpu.py -m gpu -q -x -k "$c" --timeout 120 2>&1 | tail -15 >> gpurun_out/trip1_conv.log
done
echo "=== convT" >> gpurun_out/trip1_conv.log
timeout 180 python -m pytest tests/test_conv_gpu.py -m gpu -q -k "convT" --timeout 120 2>&1 | tail -15 >> gpurun_out/trip1_conv.log
timeout 600 python -m pytest tests/test_memops_gpu.py -m gpu -q --timeout 120 2>&1 | tail -80 > gpurun_out/trip1_memops.log
grep -E "^===|passed|failed|error" gpurun_out/trip1_conv.log | head -60
tail -30 gpurun_out/trip1_memops.log
