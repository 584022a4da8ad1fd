"""Drop-in for the reference's unet/unet_parts.py (AttentionGate, DoubleConv, Down, Up, OutConv):
same names, constructor arguments and state_dict keys; forwards run sm_100a kernels."""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import ops
from ..engine import FinalConvUnit
from ..engine_unet import DoubleConvUnit, DownUnit, UpUnit
from ._base import KernelModule, from_act, to_act
from .unet_resnet import AttentionGate  # identical definition in the reference (unet_parts.py:7-30)


class DoubleConv(KernelModule):
    """(convolution => [BN] => ReLU) * 2 — unet/unet_parts.py:32-49."""

    def __init__(self, in_channels, out_channels, mid_channels=None):
        super().__init__()
        if not mid_channels:
            mid_channels = out_channels
        self.double_conv = nn.Sequential(
            nn.Conv2d(in_channels, mid_channels, kernel_size=3, padding=1, bias=False),
            nn.BatchNorm2d(mid_channels), nn.ReLU(inplace=True),
            nn.Conv2d(mid_channels, out_channels, kernel_size=3, padding=1, bias=False),
            nn.BatchNorm2d(out_channels), nn.ReLU(inplace=True))

    def _build_unit(self, sd, dev):
        return DoubleConvUnit(sd, "", dev, self.compute_dtype)

    def forward(self, x):
        self._require_eval()
        return from_act(self._unit()(x.float().contiguous()))


class Down(KernelModule):
    """Downscaling with maxpool then double conv — unet/unet_parts.py:51-63."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.maxpool_conv = nn.Sequential(nn.MaxPool2d(2), DoubleConv(in_channels, out_channels))

    def _build_unit(self, sd, dev):
        return DownUnit(sd, "", dev, self.compute_dtype)

    def forward(self, x):
        self._require_eval()
        return from_act(self._unit()(to_act(x, self.compute_dtype)))


class Up(KernelModule):
    """Upscaling then double conv — unet/unet_parts.py:65-95."""

    def __init__(self, in_channels, out_channels, bilinear=True):
        super().__init__()
        self.bilinear = bilinear
        if bilinear:
            self.up = nn.Upsample(scale_factor=2, mode='bilinear', align_corners=True)
            self.conv = DoubleConv(in_channels, out_channels, in_channels // 2)
        else:
            self.up = nn.ConvTranspose2d(in_channels, in_channels // 2, kernel_size=2, stride=2)
            self.conv = DoubleConv(in_channels, out_channels)
        self.attention = AttentionGate(F_g=in_channels // 2, F_l=in_channels // 2, F_int=in_channels // 4)

    def _build_unit(self, sd, dev):
        return UpUnit(sd, "", dev, self.compute_dtype, self.bilinear)

    def forward(self, x1, x2):
        self._require_eval()
        a1, a2 = to_act(x1, self.compute_dtype), to_act(x2, self.compute_dtype)
        psi = torch.empty((a2.N, a2.H, a2.W), dtype=torch.float32, device=a2.t.device)
        out = self._unit()(a1, a2, psi)
        self.attention.psi._emit(psi.unsqueeze(1))
        return from_act(out)


class OutConv(KernelModule):
    """unet/unet_parts.py:97-103."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.conv = nn.Conv2d(in_channels, out_channels, kernel_size=1)

    def _build_unit(self, sd, dev):
        return FinalConvUnit(sd, "conv.", dev)

    def forward(self, x):
        return self._unit()(to_act(x, self.compute_dtype))
