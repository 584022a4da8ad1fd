cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
python -m pytest tests/test_autograd_gpu.py -q -m gpu > gpurun_out/r2f_pytest_autograd.txt 2>&1; tail -5 gpurun_out/r2f_pytest_autograd.txt
for cfg in "" "VU_WGRAD_ROW3=1" "VU_WGRAD_ROW3_MIN=256" "VU_WGRAD_HALFWAVES=2" "VU_WGRAD_HALFWAVES=3" "VU_WGRAD_HALFWAVES=6" "VU_WGRAD_ROW3=1 VU_WGRAD_HALFWAVES=2"; do
  echo "== $cfg"; env $cfg python scripts/train_breakdown.py 2>&1 | tail -1
done > gpurun_out/r2f_wgrad_knobs.log 2>&1
cat gpurun_out/r2f_wgrad_knobs.log
