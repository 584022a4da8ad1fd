"""Train-mode forward of the drop-in modules through `torch.ops.vaeunet.*` (torch_ops.py), under the autograd engine.

`model.train(); seg, mu, logvar = model(images); loss.backward()` (the reference's loop, train.py:381-414) lands here for
every configuration the reference can build: UNetResNet with any `latent_injection`, `use_skip`, `use_attention`
(unet/unet_resnet.py:103-240) and the basic UNet with ConvTranspose2d or bilinear up-sampling (unet/unet_model.py:6-36,
unet/unet_parts.py:32-103). Parameters are read straight from the modules (`nn.Parameter`s, so `.grad` is populated and any
torch optimizer / GradScaler / clip_grad_norm_ works); BatchNorm running statistics and `num_batches_tracked` are updated
like `nn.BatchNorm2d` in train mode.

Activations are NHWC 16-bit tensors between the ops; `torch.cat` (pure data movement) assembles the decoder inputs and
the autograd engine sums the gradients of tensors with several consumers. Spatially constant maps (z_spatial and
everything computed from it before it meets a real feature map) stay per-image vectors: their 1x1 conv +
train-mode BatchNorm is `vec_layer`, whose statistics over the B rows equal the statistics over B*H*W identical pixels.

`vaeunet_b200.training.TrainStep` is the fused fast path for the default configuration; this path is the general one.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.nn as nn

from . import ops
from . import torch_ops  # noqa: F401  (registers torch.ops.vaeunet.*)
from ._lib import VaeUnetError

V = torch.ops.vaeunet


def _bump(bn: nn.BatchNorm2d) -> None:
    if bn.num_batches_tracked is not None:
        bn.num_batches_tracked.add_(1)


def _check_bn(bn: nn.BatchNorm2d) -> None:
    if abs(bn.eps - 1e-5) > 1e-12 or bn.momentum is None or abs(bn.momentum - 0.1) > 1e-12 or not bn.affine:
        raise VaeUnetError("train-mode BatchNorm kernels implement nn.BatchNorm2d defaults (eps 1e-5, momentum 0.1, affine)")


def bn_act(x, bn: nn.BatchNorm2d, relu: bool, residual=None):
    """train-mode BatchNorm2d (+ residual) (+ ReLU) on an NHWC 16-bit map."""
    _check_bn(bn)
    _bump(bn)
    y, mean, invstd = V.batch_norm_train(x, bn.weight, bn.bias, residual, relu)
    V.bn_update_running(bn.running_mean, bn.running_var, mean.detach(), invstd.detach(),
                        float(x.shape[0] * x.shape[1] * x.shape[2]))
    return y


def conv(x, c: nn.Conv2d):
    return V.conv2d(x, c.weight, c.bias, c.stride[0], c.padding[0])


def vec_bn(z, c: nn.Conv2d, bn: nn.BatchNorm2d, relu: bool, count: float):
    """1x1 conv + BN (+ReLU) of a spatially constant map, as a dense layer on the per-image vectors z [B, L]."""
    _check_bn(bn)
    _bump(bn)
    bias = c.bias if c.bias is not None else torch.zeros(c.weight.shape[0], device=z.device)
    y, _, mean, invstd = V.vec_layer(z, c.weight, bias, bn.weight, bn.bias, relu)
    V.bn_update_running(bn.running_mean, bn.running_var, mean.detach(), invstd.detach(), float(count))
    return y


def to_map_1ch(x, c: nn.Conv2d):
    """1x1 conv C -> 1 (+bias) -> fp32 map [N,H,W] (psi conv, final_conv, OutConv with n_classes = 1)."""
    if c.weight.shape[0] != 1:
        raise VaeUnetError("the train-mode kernels cover n_classes = 1 (the reference's lesion masks)")
    bias = c.bias if c.bias is not None else torch.zeros(1, device=x.device)
    return V.conv1x1_to_map(x, c.weight, bias)


def attention_gate(att, g, skip):
    """AttentionGate.forward in train mode (unet_resnet.py:24-29 == unet_parts.py:24-30): skip * psi."""
    g1 = bn_act(conv(g, att.W_g[0]), att.W_g[1], False)
    s = bn_act(conv(skip, att.W_x[0]), att.W_x[1], True, residual=g1)            # relu(g1 + x1)
    psi_raw = to_map_1ch(s, att.psi[0])
    bn = att.psi[1]
    _check_bn(bn)
    _bump(bn)
    psi, stat = V.map_bn_sigmoid(psi_raw, bn.weight, bn.bias)
    st = stat.detach()
    V.bn_update_running(bn.running_mean, bn.running_var, st[0:1], st[1:2], float(psi_raw.numel()))
    att.psi._emit(psi.detach().unsqueeze(1))                                   # forward hooks on .psi (analyze_model.py:733)
    return V.scale_by_map(skip, psi)


# ------------------------------------------------------------------------------------------------ UNetResNet
def encoder_features(enc, images, dt) -> List[torch.Tensor]:
    """ResNet-34 features_only in train mode (timm call site unet_resnet.py:131-137): [act1, layer1..4], NHWC 16-bit."""
    x = V.conv_small_cin(images, enc.conv1.weight, 2, 3, dt)
    f0 = bn_act(x, enc.bn1, True)
    x, _ = V.maxpool(f0, 3, 2, 1)
    feats = [f0]
    for li in range(1, 5):
        for blk in getattr(enc, f"layer{li}"):
            idt = x
            if hasattr(blk, "downsample"):
                idt = bn_act(conv(x, blk.downsample[0]), blk.downsample[1], False)
            y = bn_act(conv(x, blk.conv1), blk.bn1, True)
            x = bn_act(conv(y, blk.conv2), blk.bn2, True, residual=idt)
        feats.append(x)
    return feats


def decoder_block(blk, x, skip, z_vec, dt, x_is_vec: bool, x_hw=None):
    """DecoderBlock.forward (unet_resnet.py:71-101). x: NHWC map, or (x_is_vec) per-image vectors of a constant map of size
    x_hw. skip: NHWC map or None. z_vec: [B, L] (the constant latent map)."""
    if skip is not None:
        H, W = skip.shape[1], skip.shape[2]
    elif x_is_vec:
        H, W = x_hw[0] * 2, x_hw[1] * 2
    else:
        H, W = x.shape[1] * 2, x.shape[2] * 2
    B = z_vec.shape[0]
    if x_is_vec:
        xu = V.broadcast_vec(x, H, W, dt)               # bilinear resize of a constant map is the same constant map
    else:
        xu = V.bilinear_ac(x, H, W)
    comps = [xu]
    if skip is not None and blk.use_skip:
        comps.append(attention_gate(blk.attention, xu, skip) if blk.use_attention else skip)
    if blk.use_latent:
        zp = vec_bn(z_vec, blk.z_proj[0], blk.z_proj[1], True, B * H * W)
        comps.append(V.broadcast_vec(zp, H, W, dt))
    cat = comps[0] if len(comps) == 1 else torch.cat(comps, dim=3)
    y = bn_act(conv(cat, blk.conv1[0]), blk.conv1[1], True)
    return bn_act(conv(y, blk.conv2[0]), blk.conv2[1], True)


def unetresnet_forward(model, images: torch.Tensor, eps: Optional[torch.Tensor] = None):
    """UNetResNet.forward in train mode (unet_resnet.py:196-240) -> (logits [B,1,H,W], mu, logvar)."""
    if not images.is_cuda:
        raise VaeUnetError("input tensor is not on a CUDA device (no CPU fallback)")
    dt = model.compute_dtype
    images = images.float().contiguous()
    B, _, H, W = images.shape
    feats = encoder_features(model.encoder, images, dt)
    f4 = feats[-1]
    mu, logvar = V.latent_head(f4, model.mu_head[0].weight, model.mu_head[0].bias, model.logvar_head[0].weight,
                               model.logvar_head[0].bias)
    if model.latent_injection in ("none", "inject_no_bottleneck"):
        z = mu                                                            # unet_resnet.py:209-213: no sampling
    else:
        if eps is None:
            eps = torch.randn_like(mu)
        z = V.reparameterize(mu, logvar, eps.reshape(mu.shape).float().contiguous(), 1.0)
    h4, w4 = f4.shape[1], f4.shape[2]
    if model.use_bottleneck:
        x = vec_bn(z, model.z_initial[0], model.z_initial[1], True, B * h4 * w4)
        x_is_vec, x_hw = True, (h4, w4)
    else:
        x, x_is_vec, x_hw = f4, False, None
    for i, blk in enumerate(model.decoder_blocks):
        skip = feats[-(i + 2)] if model.use_skip else None
        x = decoder_block(blk, x, skip, z, dt, x_is_vec, x_hw)
        x_is_vec = False
    lg = to_map_1ch(x, model.final_conv)
    out = V.resize_map(lg, H, W)
    return out.unsqueeze(1), mu, logvar


# ------------------------------------------------------------------------------------------------ basic UNet
def double_conv(dc, x, dt, from_image: bool = False):
    """DoubleConv in train mode (unet_parts.py:32-49); from_image: x is the NCHW fp32 network input (Cin <= 4)."""
    seq = dc.double_conv
    if from_image:
        y = V.conv_small_cin(x, seq[0].weight, 1, 1, dt)
    else:
        y = conv(x, seq[0])
    y = bn_act(y, seq[1], True)
    return bn_act(conv(y, seq[3]), seq[4], True)


def up_block(up, x1, x2, dt):
    """Up.forward (unet_parts.py:82-95): upsample, gate the skip with the upsampled map, cat([skip_att, x_up]), DoubleConv."""
    if up.bilinear:
        xu = V.bilinear_ac(x1, x1.shape[1] * 2, x1.shape[2] * 2)
    else:
        xu = V.conv_transpose2x2(x1, up.up.weight, up.up.bias)
    if xu.shape[1] != x2.shape[1] or xu.shape[2] != x2.shape[2]:
        raise VaeUnetError("train-mode Up needs even feature-map sizes (the reference's F.pad branch is inference-only here)")
    sa = attention_gate(up.attention, xu, x2)
    return double_conv(up.conv, torch.cat([sa, xu], dim=3), dt)


def unet_forward(model, images: torch.Tensor) -> torch.Tensor:
    """UNet.forward in train mode (unet_model.py:25-36) -> logits [B,1,H,W]."""
    if not images.is_cuda:
        raise VaeUnetError("input tensor is not on a CUDA device (no CPU fallback)")
    dt = model.compute_dtype
    images = images.float().contiguous()
    if images.shape[1] > 4:
        raise VaeUnetError("the first conv runs the small-Cin kernel: n_channels <= 4")
    xs = [double_conv(model.inc, images, dt, from_image=True)]
    for d in (model.down1, model.down2, model.down3, model.down4):
        p, _ = V.maxpool(xs[-1], 2, 2, 0)
        xs.append(double_conv(d.maxpool_conv[1], p, dt))
    y = xs[4]
    for k, u in enumerate((model.up1, model.up2, model.up3, model.up4)):
        y = up_block(u, y, xs[3 - k], dt)
    lg = to_map_1ch(y, model.outc.conv)
    return lg.unsqueeze(1)
