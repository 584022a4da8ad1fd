#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests/test_training_gpu.py tests/test_train_kernels_gpu.py -m gpu -q --timeout 600 -x 2>&1 | tail -2
timeout 600 python - <<'PY'
import sys, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from vaeunet_b200 import ops, testing
from vaeunet_b200.training import TrainStep
from vaeunet_b200.unet import UNetResNet
dev = torch.device("cuda:0")
sd, kw = testing.resnet_fixture()
model = UNetResNet(**kw); model.load_state_dict(sd); model = model.to(dev).train()
ts = TrainStep(model)
ts.repack(); torch.cuda.synchronize()
# every bank against the single-tensor packers
worst = 0.0
for name, c in ts.convs.items():
    w = ts.P[c.name]
    ref = ops.pack_conv_weight(w, ts.dt)
    worst = max(worst, (c.w16.float() - ref.float()).abs().max().item())
    refb = ops.pack_dgrad_weights(w, c.stride, ts.dt)
    got = c.banks if isinstance(c.banks, (list, tuple)) else [c.banks]
    refl = refb if isinstance(refb, (list, tuple)) else [refb]
    for a, b in zip(got, refl):
        worst = max(worst, (a.float().reshape(-1) - b.float().reshape(-1)).abs().max().item())
print("batched pack vs single-tensor packers: max abs diff", worst)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): ts.repack()
e1.record(); torch.cuda.synchronize()
print("repack", e0.elapsed_time(e1) / 20, "ms")
PY
timeout 600 python scripts/train_report.py --time 16 512 2>&1 | sed -n 2,3p
} > gpurun_out/trip60.log 2>&1
cat gpurun_out/trip60.log
