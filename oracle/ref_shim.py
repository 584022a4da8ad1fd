"""Import the UNMODIFIED reference (/root/reference) in the build container (TEST INFRA).

Two `sys.modules` shims, no reference file is touched:
  * `timm` (un-vendored dependency, requirements.txt:2; call site unet/unet_resnet.py:131-137) ->
    a module whose `create_model('resnet34', features_only=True, in_chans=n)` returns a
    torchvision ResNet-34 feature net with timm's parameter names and feature order
    [act1, layer1, layer2, layer3, layer4];
  * `matplotlib` (+ .pyplot, .gridspec) -> empty modules (only plot helpers touch them).
`inject_eps` patches torch.randn_like for the duration of a call so the reference consumes
pre-generated noise. /root/reference does not exist on the GPU box: only
tests/golden/make_golden.py (run here) imports this file.
"""
from __future__ import annotations

import contextlib
import sys
import types
from pathlib import Path

import torch
import torch.nn as nn

REF = Path("/root/reference")


def available() -> bool:
    return (REF / "unet" / "unet_resnet.py").exists()


class _FeatureInfo:
    def channels(self):
        return [64, 64, 128, 256, 512]


class _ResNet34Features(nn.Module):
    def __init__(self, in_chans=3):
        super().__init__()
        import torchvision
        r = torchvision.models.resnet34(weights=None)
        self.conv1 = r.conv1 if in_chans == 3 else nn.Conv2d(in_chans, 64, 7, 2, 3, bias=False)
        self.bn1, self.act1, self.maxpool = r.bn1, nn.ReLU(inplace=True), r.maxpool
        self.layer1, self.layer2, self.layer3, self.layer4 = r.layer1, r.layer2, r.layer3, r.layer4
        self.feature_info = _FeatureInfo()

    def forward(self, x):
        x = self.act1(self.bn1(self.conv1(x)))
        feats = [x]
        x = self.maxpool(x)
        for l in (self.layer1, self.layer2, self.layer3, self.layer4):
            x = l(x)
            feats.append(x)
        return feats


def install():
    if not available():
        raise RuntimeError("/root/reference is not present (it only exists in the build container)")
    if "timm" not in sys.modules:
        timm = types.ModuleType("timm")

        def create_model(name, pretrained=False, features_only=True, in_chans=3, **kw):
            assert name == "resnet34" and features_only and not pretrained
            return _ResNet34Features(in_chans)
        timm.create_model = create_model
        sys.modules["timm"] = timm
    for n in ("matplotlib", "matplotlib.pyplot", "matplotlib.gridspec", "seaborn"):
        if n not in sys.modules:
            m = types.ModuleType(n)
            if n == "matplotlib.gridspec":
                m.GridSpec = object
            sys.modules[n] = m
    if "albumentations" not in sys.modules:   # utils/data_loading.py:9 (augmentation; never called here)
        alb = types.ModuleType("albumentations")

        def _alb_getattr(name):
            if name.startswith("__"):
                raise AttributeError(name)
            return lambda *a, **k: None
        alb.__getattr__ = _alb_getattr
        sys.modules["albumentations"] = alb
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["matplotlib"].gridspec = sys.modules["matplotlib.gridspec"]
    if str(REF) not in sys.path:
        sys.path.insert(0, str(REF))


@contextlib.contextmanager
def inject_eps(eps_list):
    """torch.randn_like(x) pops the next pre-generated tensor (must match x's shape)."""
    queue = list(eps_list)
    orig = torch.randn_like

    def fake(x, *a, **k):
        e = queue.pop(0)
        assert e.shape == x.shape, (e.shape, x.shape)
        return e.to(x.dtype)
    torch.randn_like = fake
    try:
        yield queue
    finally:
        torch.randn_like = orig
