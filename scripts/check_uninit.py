"""Debug: poison every torch.empty / empty_like allocation with NaN and look for NaN in the training step's results."""
import sys
import torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from vaeunet_b200 import ops, train_ops as T
from test_training_gpu import _setup

B, S = int(sys.argv[1]), int(sys.argv[2])
dev = torch.device("cuda:0")
O, sd, model, ts, x, m, eps = _setup("EX", dev, ops.BF16, 1.0, B, S)
ts.zero_grad()
ts.forward_backward(x.to(dev), m.to(dev), 0.001, eps.to(dev))
torch.cuda.synchronize()
g_ref = ts.flat_g.clone()

_empty, _empty_like = torch.empty, torch.empty_like
def poison(t):
    if t.is_cuda:
        if t.is_floating_point():
            t.fill_(float("nan"))
        elif t.dtype in (torch.uint8, torch.int8):
            t.fill_(127)
    return t
torch.empty = lambda *a, **k: poison(_empty(*a, **k))
torch.empty_like = lambda *a, **k: poison(_empty_like(*a, **k))
ts.zero_grad()
out = ts.forward_backward(x.to(dev), m.to(dev), 0.001, eps.to(dev))
torch.cuda.synchronize()
torch.empty, torch.empty_like = _empty, _empty_like
print("loss", out["loss"].item())
bad = [(k, int(torch.isnan(g).sum().item()), g.numel()) for k, g in ts.G.items() if torch.isnan(g).any()]
print(f"{len(bad)} gradient tensors with NaN of {len(ts.G)}")
for k, n, tot in bad[:60]:
    print(f"  {k}: {n}/{tot}")
g = ts.flat_g
ok = ~torch.isnan(g)
print("rel diff on the finite part vs unpoisoned run:", ((g[ok] - g_ref[ok]).norm() / g_ref[ok].norm()).item())
