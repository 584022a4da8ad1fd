"""Deterministic fixture weights (seeded init + committed BN-calibration overlay) used by the
parity tests, `__graft_entry__.smoke()` and `bench.py` (there are no shipped checkpoints)."""
from __future__ import annotations

import math
from pathlib import Path
from typing import Dict

import torch

GOLDEN = Path(__file__).resolve().parent.parent / "tests" / "golden"
# BN gamma of the last BatchNorm of every residual block. timm zero-initialises it (zero_init_last) and
# trained ResNets keep it small; 0.25 keeps the seeded network from amplifying rounding noise the way
# a fully random residual stack does (DESIGN.md "Fixtures").
RESIDUAL_GAMMA = 0.25


def seeded_state_dict(template: Dict[str, torch.Tensor], seed: int = 0) -> Dict[str, torch.Tensor]:
    """Fill a state_dict template (keys/shapes) with seeded values, in key order, on the CPU."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for k, t in template.items():
        shp = tuple(t.shape)
        if k.endswith("num_batches_tracked"):
            v = torch.zeros(shp, dtype=torch.int64)
        elif k.endswith("running_mean"):
            v = torch.zeros(shp)
        elif k.endswith("running_var"):
            v = torch.ones(shp)
        elif len(shp) == 4:                              # conv / conv-transpose weight
            if ".up.weight" in k:                        # ConvTranspose2d [I,O,2,2]
                fan_in = shp[0]
            else:
                fan_in = shp[1] * shp[2] * shp[3]
            v = torch.randn(shp, generator=g) * math.sqrt(2.0 / fan_in)
        elif k.endswith(".weight"):                      # BN gamma
            base = RESIDUAL_GAMMA if (".bn2." in k) else 1.0
            v = base + 0.1 * torch.randn(shp, generator=g)
        elif k.endswith(".bias"):
            v = 0.1 * torch.randn(shp, generator=g)
        else:
            raise KeyError(k)
        sd[k] = v
    return sd


def load_fixture(name: str, template: Dict[str, torch.Tensor], seed: int = 0) -> Dict[str, torch.Tensor]:
    """seeded init + overlay of the small calibrated tensors in tests/golden/<name>.pt
    (BN running stats, rescaled heads / output conv) written by tests/golden/make_golden.py."""
    sd = seeded_state_dict(template, seed)
    f = GOLDEN / f"{name}.pt"
    if not f.exists():
        raise FileNotFoundError(f"{f} missing: run tests/golden/make_golden.py in the build container")
    overlay = torch.load(f, map_location="cpu", weights_only=True)
    for k, v in overlay.items():
        assert k in sd and sd[k].shape == v.shape, k
        sd[k] = v.clone()
    return sd


def resnet_fixture(seed: int = 0, **model_kwargs):
    """-> (state_dict on CPU, model kwargs). Template comes from the drop-in module itself."""
    from .unet.unet_resnet import UNetResNet
    kw = dict(n_channels=3, n_classes=1, pretrained=False)
    kw.update(model_kwargs)
    tmpl = UNetResNet(**kw).state_dict()
    name = "fixture_resnet" + ("" if kw.get("latent_injection", "all") == "all" else "_" + str(kw["latent_injection"]))
    return load_fixture(name, tmpl, seed), kw


def unet_fixture(seed: int = 0, bilinear: bool = False):
    from .unet.unet_model import UNet
    tmpl = UNet(3, 1, bilinear).state_dict()
    return load_fixture("fixture_unet" + ("_bilinear" if bilinear else ""), tmpl, seed)


def resnet_variant_state(seed: int = 0, **model_kwargs):
    """State for a UNetResNet VARIANT without its own calibrated fixture (use_skip=False, use_attention=False, ...): seeded
    init overlaid with every tensor of the default fixture whose key and shape match (encoder statistics, rescaled heads,
    output conv). -> (state_dict on CPU, model kwargs)."""
    from .unet.unet_resnet import UNetResNet
    kw = dict(n_channels=3, n_classes=1, pretrained=False)
    kw.update(model_kwargs)
    tmpl = UNetResNet(**kw).state_dict()
    sd = seeded_state_dict(tmpl, seed)
    overlay = torch.load(GOLDEN / "fixture_resnet.pt", map_location="cpu", weights_only=True)
    for k, v in overlay.items():
        if k in sd and sd[k].shape == v.shape and not k.startswith("decoder_blocks."):
            sd[k] = v.clone()
    return sd, kw


def f3_inputs(seed: int, B: int, H: int, W: int, ties: bool):
    """Correlated synthetic (probabilities, ground truth, uncertainty): uncertainty is high where the prediction is wrong."""
    g = torch.Generator().manual_seed(seed)
    gt = (torch.rand(B, H, W, generator=g) > 0.8).float()
    logit = (gt * 2 - 1) * 1.5 + torch.randn(B, H, W, generator=g) * 1.8
    p = torch.sigmoid(logit)
    wrong = ((p > 0.5).float() != gt).float()
    unc = (0.05 + 0.10 * wrong + 0.35 * torch.rand(B, H, W, generator=g) + 0.3 * (p * (1 - p))).clamp(0, 1)
    if ties:
        unc = torch.round(unc * 50) / 50          # heavy ties: sklearn groups equal thresholds
    return p, gt, unc


def f4_inputs(seed: int, H: int, W: int):
    """A synthetic 8-bit 'decoded fundus photograph' [H,W,3] with a black border outside the disc, and an 8-bit lesion mask
    [H,W] (0 / 255) - the inputs of the patch pipeline tests."""
    import numpy as np
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:H, 0:W]
    r = np.hypot((yy - H / 2) / (0.5 * H), (xx - W / 2) / (0.5 * W))
    disc = r < 0.92
    img = np.zeros((H, W, 3), dtype=np.float64)
    base = np.stack([180 - 60 * r, 90 - 40 * r, 40 - 10 * r], -1)
    img[disc] = (base + rng.normal(0, 12, (H, W, 3)))[disc]
    img = np.clip(img, 0, 255).astype(np.uint8)
    mask = np.zeros((H, W), dtype=np.uint8)
    for _ in range(6):
        cy, cx, rad = rng.integers(H // 6, 5 * H // 6), rng.integers(W // 6, 5 * W // 6), rng.integers(3, max(4, H // 20))
        mask[np.hypot(yy - cy, xx - cx) < rad] = 255
    return img, mask
