"""Per-tensor gradient cosines of the basic-UNet train-mode autograd path vs the fp32 oracle (debug aid)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from oracle import vaeunet_oracle as O
from vaeunet_b200 import testing
from vaeunet_b200.synthetic import synthetic_fundus
from vaeunet_b200.unet import UNet
from vaeunet_b200.utils.loss import CombinedLoss

dev = torch.device("cuda:0")
SCALE = float(sys.argv[2]) if len(sys.argv) > 2 else 1024.0
dtype = sys.argv[3] if len(sys.argv) > 3 else "f16"
for bilinear in (False, True):
    sd = testing.unet_fixture(bilinear=bilinear)
    model = UNet(3, 1, bilinear); model.load_state_dict(sd); model = model.to(dev); model.compute_dtype = dtype
    B, S = 4, int(sys.argv[1]) if len(sys.argv) > 1 else 128
    x, m = synthetic_fundus(B, S, S, seed=51)
    leaves = {k: (v.detach().clone().requires_grad_(True) if v.is_floating_point() and not k.endswith(("running_mean", "running_var")) else v)
              for k, v in sd.items()}
    st = O.BNState(training=True)
    out = O.unet_forward(leaves, x, bilinear, st)
    loss = O.combined_loss(out, m)
    names = [k for k, v in leaves.items() if isinstance(v, torch.Tensor) and v.requires_grad]
    gs = torch.autograd.grad(loss, [leaves[k] for k in names], allow_unused=True)
    gr = {k: (g if g is not None else torch.zeros_like(leaves[k])) for k, g in zip(names, gs)}
    model.train(); model.zero_grad(set_to_none=True)
    o = model(x.to(dev))
    l = CombinedLoss()(o, m.to(dev))
    (l * SCALE).backward()
    torch.cuda.synchronize()
    print(f"bilinear={bilinear} loss {l.item():.6f} vs {loss.item():.6f}; logits rel {((o.cpu()-out).norm()/out.norm()).item():.2e}")
    cos = lambda a, b: (a.double().flatten() @ b.double().flatten() / (a.double().norm() * b.double().norm() + 1e-300)).item()
    for k, p in model.named_parameters():
        g = p.grad.cpu() / SCALE
        if p.dim() == 4 or "up.bias" in k:
            print(f"  {k:50s} cos {cos(g, gr[k]):.4f}  |g| {g.norm().item():.3e} |ref| {gr[k].norm().item():.3e}  max|g*SCALE| {p.grad.abs().max().item():.3e}")
