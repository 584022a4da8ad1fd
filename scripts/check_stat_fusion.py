"""Debug: per-BatchNorm (sums, mean, invstd, output) of the training forward compared between repeated identical runs."""
import sys
import torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from vaeunet_b200 import ops, train_ops as T
from test_training_gpu import _setup

B, S = int(sys.argv[1]), int(sys.argv[2])
dev = torch.device("cuda:0")
O, sd, model, ts, x, m, eps = _setup("EX", dev, ops.BF16, 1.0, B, S)
orig = T.bn_train_fwd
rec = []

def patched(xa, *a, sums=None, **k):
    y, mean, invstd = orig(xa, *a, sums=sums, **k)
    rec.append(dict(name=f"N{xa.N} {xa.H}x{xa.W} C{xa.C} fused={sums is not None}", x=xa.t.clone(), y=y.t.clone(), mean=mean.clone(),
                    invstd=invstd.clone(), sums=None if sums is None else sums.clone()))
    return y, mean, invstd

T.bn_train_fwd = patched
runs, gs = [], []
for run in range(3):
    rec.clear()
    ts.zero_grad()
    out = ts.forward_backward(x.to(dev), m.to(dev), 0.001, eps.to(dev))
    torch.cuda.synchronize()
    gs.append(ts.flat_g.clone())
    runs.append(list(rec))
    print(f"run {run}: loss {out['loss'].item():.9f}")
r0, r1 = runs[0], runs[1]
rel = lambda a, b: ((a.double() - b.double()).norm() / (b.double().norm() + 1e-30)).item()
for i, (a, b) in enumerate(zip(r0, r1)):
    line = f"{i:3d} {a['name']:36s} x {rel(a['x'], b['x']):.2e} mean {rel(a['mean'], b['mean']):.2e} invstd {rel(a['invstd'], b['invstd']):.2e} y {rel(a['y'], b['y']):.2e}"
    if a["sums"] is not None:
        line += f" sums {rel(a['sums'], b['sums']):.2e}"
    print(line)
print("run-to-run rel diff of the gradient:", [((gs[0] - g).norm() / gs[0].norm()).item() for g in gs[1:]])
