"""Seeded synthetic fundus patches: dark background, an orange disc with a smooth
radial gradient and low-frequency noise, bright Gaussian blobs (EX-like) or tiny dark-red blobs
(MA-like); mask = thresholded blob field. CPU generator => identical on every machine."""
from __future__ import annotations

import torch
import torch.nn.functional as F


def synthetic_fundus(batch: int, height: int, width: int = None, seed: int = 1234, lesion: str = "EX"):
    """-> images fp32 [B,3,H,W] in [0,1], masks fp32 [B,1,H,W] in {0,1}."""
    width = width or height
    g = torch.Generator().manual_seed(seed)
    H, W = height, width
    S = float(min(H, W))
    yy = torch.arange(H, dtype=torch.float32).view(1, H, 1)
    xx = torch.arange(W, dtype=torch.float32).view(1, 1, W)
    cy = H / 2 + (torch.rand(batch, 1, 1, generator=g) - 0.5) * 0.1 * H
    cx = W / 2 + (torch.rand(batch, 1, 1, generator=g) - 0.5) * 0.1 * W
    r = torch.sqrt((yy - cy) ** 2 + (xx - cx) ** 2) / (0.48 * max(H, W))
    disc = torch.sigmoid((1.0 - r) * 40.0)
    shade = (1.0 - 0.45 * r.clamp(max=1.0) ** 2)
    lo = torch.randn(batch, 3, max(H // 32, 2), max(W // 32, 2), generator=g) * 0.05
    lo = F.interpolate(lo, size=(H, W), mode="bilinear", align_corners=True)
    base = torch.tensor([0.78, 0.42, 0.16]).view(1, 3, 1, 1)
    img = (base * shade.unsqueeze(1) + lo) * disc.unsqueeze(1) + 0.02
    field = torch.zeros(batch, H, W)
    for b in range(batch):
        if lesion == "MA":
            n = int(torch.randint(20, 61, (1,), generator=g))
            sig = 1.0 + torch.rand(n, generator=g)
        else:
            n = int(torch.randint(5, 21, (1,), generator=g))
            sig = (2.0 + 6.0 * torch.rand(n, generator=g)) * (S / 512.0) + 0.5
        by = (0.15 + 0.7 * torch.rand(n, generator=g)) * H
        bx = (0.15 + 0.7 * torch.rand(n, generator=g)) * W
        amp = 0.6 + 0.4 * torch.rand(n, generator=g)
        for k in range(n):
            field[b] += amp[k] * torch.exp(-((yy[0] - by[k]) ** 2 + (xx[0] - bx[k]) ** 2) / (2 * sig[k] ** 2))
    field = field * disc
    if lesion == "MA":
        img = img - field.unsqueeze(1) * torch.tensor([0.35, 0.30, 0.10]).view(1, 3, 1, 1)
    else:
        img = img + field.unsqueeze(1) * torch.tensor([0.25, 0.45, 0.35]).view(1, 3, 1, 1)
    img = img.clamp(0.0, 1.0).contiguous()
    mask = (field > 0.35).float().unsqueeze(1).contiguous()
    return img, mask
