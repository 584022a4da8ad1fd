"""Debug: is one decode's result independent of how many other decodes share the launch? (sample-sharded MC relies on it)"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from vaeunet_b200 import testing
from vaeunet_b200.synthetic import synthetic_fundus
from vaeunet_b200.unet import UNetResNet

dev = torch.device("cuda:0")
sd, kw = testing.resnet_fixture()
model = UNetResNet(**kw); model.load_state_dict(sd); model = model.to(dev).eval()
for S in (700, 512):
    x, _ = synthetic_fundus(1, S, seed=21)
    x = x.to(dev)
    eps = torch.randn(15, 1, 32, generator=torch.Generator().manual_seed(3)).to(dev)
    eng = model.engine()
    xf = x.float().contiguous()
    full, _, _ = eng.mc_logits(xf, 2.0, 15, eps, True)
    a, _, _ = eng.mc_logits(xf, 2.0, 8, eps[:8], True)
    b, _, _ = eng.mc_logits(xf, 2.0, 7, eps[8:], True)
    ps = torch.cat([a, b], dim=1)
    d = (full.reshape(15, -1) - ps.reshape(15, -1))
    print(S, "logit scale", full.abs().max().item(), "per-sample max-abs diff of the logits:", [f"{v:.1e}" for v in d.abs().max(dim=1).values.tolist()])
