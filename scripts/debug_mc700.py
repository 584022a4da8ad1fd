"""Parity margins of mc_predict at 700x700, T=2, N=15 (configs[3]) against the fp32 oracle, per output map.
Run under different VU_* switches to attribute a change: VU_NO_FAST_CONCAT=1, VU_STEM_FMA=1."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from oracle import vaeunet_oracle as O
from vaeunet_b200 import testing
from vaeunet_b200.synthetic import synthetic_fundus
from vaeunet_b200.unet import UNetResNet

dev = torch.device("cuda:0")
sd, kw = testing.resnet_fixture()
m = UNetResNet(**kw); m.load_state_dict(sd); m = m.to(dev).eval(); m.compute_dtype = "f16"
cfg = O.ResNetCfg()
rel = lambda a, b: ((a.float().cpu() - b.float().cpu()).norm() / (b.float().cpu().norm() + 1e-12)).item()
for (S, T, N, seed) in ((700, 2.0, 15, 703), (512, 1.0, 10, 11)):
    x, _ = synthetic_fundus(1, S, S, seed=seed, lesion="MA")
    eps = torch.randn(N, 1, 32, generator=torch.Generator().manual_seed(seed + 1))
    ref = O.mc_patch_uncertainty(sd, cfg, x, T, N, eps)
    with torch.no_grad():
        res = m.mc_predict(x.to(dev), T, N, eps=eps.to(dev))
    print(f"S={S} T={T} N={N} env: fastcat={'VU_NO_FAST_CONCAT' not in os.environ} stem_mma={'VU_STEM_FMA' not in os.environ}: "
          + " ".join(f"{k}={rel(res[k], ref[k]):.2e}" for k in ("mean", "std", "entropy", "mutual_info", "mu", "logvar")))
