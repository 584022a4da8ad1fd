"""Is the training step launch-bound? Host time to ENQUEUE one step (no sync) beside the device time of the step."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from vaeunet_b200 import ops, testing
from vaeunet_b200.synthetic import synthetic_fundus
from vaeunet_b200.training import TrainStep
from vaeunet_b200.unet import UNetResNet

dev = torch.device("cuda:0")
sd, kw = testing.resnet_fixture()
model = UNetResNet(**kw); model.load_state_dict(sd); model = model.to(dev).train()
ts = TrainStep(model, lesion="EX", dtype=ops.BF16)
B = 16
x, m = synthetic_fundus(4, 512, seed=1)
x, m = x.repeat(B // 4, 1, 1, 1).to(dev), m.repeat(B // 4, 1, 1, 1).to(dev)
eps = torch.randn(B, 32, device=dev)
for _ in range(3):
    ts.step(x, m, 0.001, eps)
torch.cuda.synchronize()
K = 10
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
host = []
e0.record()
for _ in range(K):
    t = time.perf_counter()
    ts.step(x, m, 0.001, eps)
    host.append((time.perf_counter() - t) * 1e3)
t_enq = sum(host)
e1.record()
torch.cuda.synchronize()
print(f"device {e0.elapsed_time(e1) / K:.2f} ms/step; host enqueue {t_enq / K:.2f} ms/step (min {min(host):.2f}, max {max(host):.2f})")
# one step enqueued from an idle GPU with a sync before: the host runs ahead only if enqueue < device
torch.cuda.synchronize()
t = time.perf_counter(); ts.step(x, m, 0.001, eps); te = time.perf_counter() - t; torch.cuda.synchronize(); tt = time.perf_counter() - t
print(f"single step from idle: enqueue {te * 1e3:.2f} ms, complete {tt * 1e3:.2f} ms")
