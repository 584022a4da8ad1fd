"""Debug: which gradient tensors differ between two identical training steps."""
import sys
import torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from vaeunet_b200 import ops
from test_training_gpu import _setup

B, S = int(sys.argv[1]), int(sys.argv[2])
dev = torch.device("cuda:0")
O, sd, model, ts, x, m, eps = _setup("EX", dev, ops.BF16, 1.0, B, S)
gs = []
for run in range(3):
    ts.zero_grad()
    ts.forward_backward(x.to(dev), m.to(dev), 0.001, eps.to(dev))
    torch.cuda.synchronize()
    gs.append({k: g.clone() for k, g in ts.G.items()})
tot = torch.sqrt(sum((g.double() ** 2).sum() for g in gs[0].values())).item()
for other in (1, 2):
    rows = sorted(((gs[0][k] - gs[other][k]).norm().item(), k) for k in gs[0])[::-1][:8]
    print(f"run 0 vs {other}: total norm {tot:.4e}")
    for d, k in rows:
        print(f"   |diff| {d:.4e}  |g| {gs[0][k].norm().item():.4e}  {k}")
