"""In-situ kernel times of the training step (B=16, 512x512) from the CUPTI activity trace (torch.profiler): unlike the ncu
launch list these are warm-cache, un-serialised durations with the weight-gradient stream overlapping the main chain, so
the per-family sums show where the 13 ms of the real step go. usage: train_kineto.py [B] [S] [steps]"""
import sys
from collections import defaultdict
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from torch.profiler import ProfilerActivity, profile
from vaeunet_b200 import testing
from vaeunet_b200.synthetic import synthetic_fundus
from vaeunet_b200.training import TrainStep
from vaeunet_b200.unet import UNetResNet

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
S = int(sys.argv[2]) if len(sys.argv) > 2 else 512
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda:0")
sd, kw = testing.resnet_fixture()
model = UNetResNet(**kw); model.load_state_dict(sd); model = model.to(dev).train()
ts = TrainStep(model)
x, m = synthetic_fundus(2, S, S, seed=5)
x = x.repeat(B // 2, 1, 1, 1).to(dev); m = m.repeat(B // 2, 1, 1, 1).to(dev)
eps = torch.randn(B, 32, device=dev)
for _ in range(3):
    ts.step(x, m, 0.001, eps)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(steps):
        ts.step(x, m, 0.001, eps)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
t0 = min(e.time_range.start for e in ev); t1 = max(e.time_range.end for e in ev)
agg = defaultdict(lambda: [0.0, 0])
for e in ev:
    name = e.name.split("(")[0].replace("void ", "").replace("vu::", "")
    agg[name][0] += (e.time_range.end - e.time_range.start) / steps
    agg[name][1] += 1
tot = sum(v[0] for v in agg.values())
print(f"span per step {(t1 - t0) / steps / 1e3:.3f} ms; sum of kernel durations per step {tot / 1e3:.3f} ms (streams overlap)")
for name, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:45]:
    print(f"{us / 1e3:8.3f} ms {100 * us / tot:5.1f}%  x{n // steps:4d}  {name[:90]}")
