"""Packed-weight units of the basic attention U-Net (unet/unet_parts.py, unet/unet_model.py)."""
from __future__ import annotations

from typing import Optional, Union

import torch

from . import ops
from .engine import ConvUnit, FinalConvUnit, GateUnit, fold_bn
from .ops import Act


class DoubleConvUnit:
    """(conv3x3 no-bias -> BN -> ReLU) x 2, unet/unet_parts.py:32-49. A first conv with Cin <= 4
    (the network input) runs the direct small-Cin kernel from NCHW fp32."""

    def __init__(self, sd, prefix: str, dev, dtype):
        p = prefix + "double_conv."
        self.dtype = dtype
        w0 = sd[p + "0.weight"]
        self.small = w0.shape[1] <= 4
        if self.small:
            self.w0 = w0.detach().to(dev, torch.float32).contiguous()
            self.s0, self.b0 = fold_bn(sd, p + "1.", None, dev)
        else:
            self.c0 = ConvUnit(sd, p + "0.", p + "1.", 3, 1, 1, dev, dtype)
        self.c1 = ConvUnit(sd, p + "3.", p + "4.", 3, 1, 1, dev, dtype)

    def __call__(self, x: Union[Act, torch.Tensor]) -> Act:
        if self.small:
            if isinstance(x, Act):
                x = ops.nhwc_to_nchw(x)
            y = ops.conv3x3_small_cin(x, self.w0, self.s0, self.b0, True, self.dtype)
        else:
            if not isinstance(x, Act):
                x = ops.nchw_to_nhwc(x, dtype=self.dtype)
            y = self.c0(x, relu=True)
        return self.c1(y, relu=True)


class DownUnit:
    """MaxPool2d(2) -> DoubleConv, unet/unet_parts.py:51-63."""

    def __init__(self, sd, prefix: str, dev, dtype):
        self.dc = DoubleConvUnit(sd, prefix + "maxpool_conv.1.", dev, dtype)

    def __call__(self, x: Act) -> Act:
        return self.dc(ops.maxpool(x, 2, 2, 0))


class UpUnit:
    """Up.forward, unet/unet_parts.py:82-95: upsample (ConvTranspose2d k2 s2 as four tensor-core
    1x1 GEMMs with a pixel-shuffle store, or bilinear x2), pad to the skip, gate the skip,
    cat([skip_att, x_up]) by slice writes, DoubleConv."""

    def __init__(self, sd, prefix: str, dev, dtype, bilinear: bool):
        self.bilinear = bilinear
        self.dtype = dtype
        if not bilinear:
            w = sd[prefix + "up.weight"].detach().to(dev, torch.float32).contiguous()   # [I][O][2][2]
            self.up_w = ops.pack_convT2x2_weight(w, dtype)
            self.up_b = sd[prefix + "up.bias"].detach().to(dev, torch.float32).contiguous()
            self.up_cout = w.shape[1]
        self.gate = GateUnit(sd, prefix + "attention.", dev, dtype)
        self.dc = DoubleConvUnit(sd, prefix + "conv.", dev, dtype)

    def __call__(self, x1: Act, x2: Act, psi_out=None) -> Act:
        cu = x1.C if self.bilinear else self.up_cout
        cs = x2.C
        cat = ops.new_act(x2.N, x2.H, x2.W, cs + cu, self.dtype, x2.t.device)
        up = cat.slice(cs, cu)
        Hu, Wu = x1.H * 2, x1.W * 2
        dy, dx = x2.H - Hu, x2.W - Wu
        if dy < 0 or dx < 0:
            raise ValueError("Up: upsampled map larger than the skip connection")
        padded = (dy != 0 or dx != 0)
        if padded:
            cat.t[..., cat.coff + cs: cat.coff + cs + cu].zero_()
        if self.bilinear:
            ops.bilinear_ac(x1, Hu, Wu, out=up, pad_t=dy // 2, pad_l=dx // 2)
        else:
            tgt = up
            if padded:
                tgt = ops.new_act(x1.N, Hu, Wu, cu, self.dtype, x1.t.device)
            for ab in range(4):
                ops.conv2d(x1, self.up_w[ab], 1, 1, 0, None, self.up_b, False, out=tgt, out_up=2, out_dh=ab // 2,
                           out_dw=ab % 2)
            if padded:
                ops.bilinear_ac(tgt, Hu, Wu, out=up, pad_t=dy // 2, pad_l=dx // 2)  # identity resize = padded copy
        self.gate(up, x2, cat.slice(0, cs), psi_out)
        return self.dc(cat)


class UNetEngine:
    """UNet.forward, unet/unet_model.py:25-36."""

    def __init__(self, sd, device, bilinear=False, dtype=ops.DEFAULT_DTYPE):
        ops.require_device()
        dev = torch.device(device)
        self.dev, self.dtype = dev, dtype
        self.inc = DoubleConvUnit(sd, "inc.", dev, dtype)
        self.downs = [DownUnit(sd, f"down{k}.", dev, dtype) for k in range(1, 5)]
        self.ups = [UpUnit(sd, f"up{k}.", dev, dtype, bilinear) for k in range(1, 5)]
        self.outc = FinalConvUnit(sd, "outc.conv.", dev)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        xs = [self.inc(x.float().contiguous())]
        for d in self.downs:
            xs.append(d(xs[-1]))
        y = xs[4]
        for k, u in enumerate(self.ups):
            y = u(y, xs[3 - k])
        return self.outc(y)
