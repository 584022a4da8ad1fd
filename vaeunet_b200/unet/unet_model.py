"""Drop-in for the reference's unet/unet_model.py — UNet(n_channels, n_classes, bilinear=False)."""
from __future__ import annotations

import torch.nn as nn

from ..engine_unet import UNetEngine
from ._base import KernelModule
from .unet_parts import DoubleConv, Down, OutConv, Up


class UNet(KernelModule):
    """unet/unet_model.py:6-36: inc 64 -> down 128/256/512/1024(/factor) -> up x4 (attention-gated
    skips) -> outc. forward returns logits [B, n_classes, H, W]."""

    def __init__(self, n_channels, n_classes, bilinear=False):
        super().__init__()
        self.n_channels = n_channels
        self.n_classes = n_classes
        self.bilinear = bilinear
        self.inc = DoubleConv(n_channels, 64)
        self.down1 = Down(64, 128)
        self.down2 = Down(128, 256)
        self.down3 = Down(256, 512)
        factor = 2 if bilinear else 1
        self.down4 = Down(512, 1024 // factor)
        self.up1 = Up(1024, 512 // factor, bilinear)
        self.up2 = Up(512, 256 // factor, bilinear)
        self.up3 = Up(256, 128 // factor, bilinear)
        self.up4 = Up(128, 64, bilinear)
        self.outc = OutConv(64, n_classes)

    def _build_unit(self, sd, dev):
        return UNetEngine(sd, dev, self.bilinear, self.compute_dtype)

    def forward(self, x):
        if self.training:
            # train mode (batch-statistics BatchNorm, autograd): torch.ops.vaeunet.* custom ops
            from ..autograd_path import unet_forward
            return unet_forward(self, x)
        return self._unit().forward(x)

    def use_checkpointing(self):
        """The reference's version (unet_model.py:38-48) calls the *module* torch.utils.checkpoint and
        raises TypeError; inference needs no activation checkpointing, so this is a documented no-op."""
        return self
