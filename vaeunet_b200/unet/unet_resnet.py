"""Drop-in for the reference's unet/unet_resnet.py: AttentionGate, DecoderBlock, UNetResNet.

Same class names, constructor arguments, sub-module attribute names, `state_dict` keys and return
types as the reference (unet/unet_resnet.py:6-279), so the reference's scripts that reach inside
the model (`model.encoder(x)`, `model.mu_head`, `model.decoder_blocks[i](x, skip, z)`,
`model.final_conv`, utils/vae_utils.py:18-67, visualize_vae.py:68-79) keep working. Every
forward runs hand-written sm_100a kernels; nothing falls back to torch ops.

The reference builds its encoder with `timm.create_model('resnet34', features_only=True)`
(unet_resnet.py:131-137; timm is an un-vendored dependency). `ResNet34Features` below declares the
same parameter names (conv1, bn1, layer{1-4}.{i}.{conv1,bn1,conv2,bn2}, downsample.{0,1}) so timm /
torchvision ResNet-34 checkpoints load unchanged; `pretrained=True` needs a network and raises.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.nn as nn

from .. import ops
from ..engine import (ConstVec, DecoderBlockUnit, EncoderUnit, FinalConvUnit, GateUnit, HeadUnit,
                      ResNetVAEEngine, ZInitialUnit, _latent_flags)
from ._base import KernelModule, from_act, to_act


class AttentionGate(KernelModule):
    """unet/unet_resnet.py:6-29."""

    def __init__(self, F_g, F_l, F_int):
        super().__init__()
        self.W_g = nn.Sequential(nn.Conv2d(F_g, F_int, kernel_size=1), nn.BatchNorm2d(F_int))
        self.W_x = nn.Sequential(nn.Conv2d(F_l, F_int, kernel_size=1), nn.BatchNorm2d(F_int))
        self.psi = _PsiHead(F_int)
        self.relu = nn.ReLU(inplace=True)

    def _build_unit(self, sd, dev):
        return GateUnit(sd, "", dev, self.compute_dtype)

    def forward(self, g, x):
        self._require_eval()
        unit = self._unit()
        ga, xa = to_act(g, self.compute_dtype), to_act(x, self.compute_dtype)
        out = ops.new_act(ga.N, xa.H, xa.W, xa.C, ga.dt, ga.t.device)
        psi = torch.empty((ga.N, xa.H, xa.W), dtype=torch.float32, device=ga.t.device)
        unit(ga, xa, out, psi)
        self.psi._emit(psi.unsqueeze(1))  # forward hooks on .psi (analyze_model.py:733-738)
        return from_act(out)


class _PsiHead(nn.Sequential):
    """Container for psi = Conv1x1(F_int->1) + BN(1) + Sigmoid. The arithmetic happens inside the
    fused gate kernel; `_emit` replays the result through the module so forward hooks fire."""

    def __init__(self, F_int):
        super().__init__(nn.Conv2d(F_int, 1, kernel_size=1), nn.BatchNorm2d(1), nn.Sigmoid())

    def forward(self, psi_map):
        return psi_map

    def _emit(self, psi_map):
        return self(psi_map)


class DecoderBlock(KernelModule):
    """unet/unet_resnet.py:31-101."""

    def __init__(self, in_channels, skip_channels, out_channels, latent_dim, use_attention=True, use_skip=True,
                 use_latent=True):
        super().__init__()
        self.use_latent = use_latent
        self.latent_dim = latent_dim
        if use_latent:
            self.z_proj = nn.Sequential(nn.Conv2d(latent_dim, latent_dim, 1), nn.BatchNorm2d(latent_dim),
                                        nn.ReLU(inplace=True))
        self.use_skip = use_skip
        self.use_attention = use_attention and use_skip
        if self.use_attention:
            self.attention = AttentionGate(in_channels, skip_channels, in_channels // 4)
        input_channels = in_channels
        if use_skip:
            input_channels += skip_channels
        if use_latent:
            input_channels += latent_dim
        self.conv1 = nn.Sequential(nn.Conv2d(input_channels, out_channels, 3, padding=1, bias=False),
                                   nn.BatchNorm2d(out_channels), nn.ReLU(inplace=True))
        self.conv2 = nn.Sequential(nn.Conv2d(out_channels, out_channels, 3, padding=1, bias=False),
                                   nn.BatchNorm2d(out_channels), nn.ReLU(inplace=True))

    def _build_unit(self, sd, dev):
        return DecoderBlockUnit(sd, "", dev, self.compute_dtype, self.latent_dim, self.use_attention, self.use_skip,
                                self.use_latent)

    def forward(self, x, skip, z):
        self._require_eval()
        unit = self._unit()
        xa = to_act(x, self.compute_dtype)
        sa = to_act(skip, self.compute_dtype) if skip is not None else None
        za = None
        if self.use_latent:
            za = ConstVec(z.float().reshape(z.shape[0], z.shape[1]).contiguous()) if z.shape[2:] == (1, 1) else to_act(z, self.compute_dtype)
        psi = None
        if self.use_attention and sa is not None:
            psi = torch.empty((xa.N, sa.H, sa.W), dtype=torch.float32, device=xa.t.device)
        out = unit(xa, sa, za, psi)
        if psi is not None:
            self.attention.psi._emit(psi.unsqueeze(1))
        return from_act(out)


class _BasicBlock(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)
        if stride != 1 or cin != cout:
            self.downsample = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))


class _FeatureInfo:
    def __init__(self, ch):
        self._ch = ch

    def channels(self):
        return list(self._ch)


class ResNet34Features(KernelModule):
    """Stand-in for `timm.create_model('resnet34', features_only=True, in_chans=n)`:
    callable -> [act1, layer1, layer2, layer3, layer4] (channels 64,64,128,256,512; strides 2..32)."""

    def __init__(self, in_chans=3):
        super().__init__()
        self.conv1 = nn.Conv2d(in_chans, 64, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        cfg = [(64, 64, 3, 1), (64, 128, 4, 2), (128, 256, 6, 2), (256, 512, 3, 2)]
        for li, (cin, cout, n, stride) in enumerate(cfg, start=1):
            blocks = [_BasicBlock(cin if b == 0 else cout, cout, stride if b == 0 else 1) for b in range(n)]
            setattr(self, f"layer{li}", nn.Sequential(*blocks))
        self.feature_info = _FeatureInfo([64, 64, 128, 256, 512])

    def _build_unit(self, sd, dev):
        return EncoderUnit(sd, "", dev, self.compute_dtype)

    def forward_acts(self, x) -> List[ops.Act]:
        self._require_eval()
        if not x.is_cuda:
            from .._lib import VaeUnetError
            raise VaeUnetError("input tensor is not on a CUDA device (no CPU fallback)")
        return self._unit()(x.float().contiguous())

    def forward(self, x):
        return [from_act(a) for a in self.forward_acts(x)]


class _LatentHead(KernelModule, nn.Sequential):
    """mu_head / logvar_head = Conv1x1(512->L) + AdaptiveAvgPool2d(1) (unet_resnet.py:140-147).
    Callable on NCHW features -> [B, L, 1, 1]."""

    def __init__(self, cin, latent_dim):
        nn.Sequential.__init__(self, nn.Conv2d(cin, latent_dim, kernel_size=1), nn.AdaptiveAvgPool2d(1))

    def _build_unit(self, sd, dev):
        return HeadUnit(sd, "", dev)

    def forward(self, x_enc):
        u = self._unit()
        a = to_act(x_enc, self.compute_dtype)
        mu, _, _ = ops.latent_head(a, u.w, u.b, u.w, u.b, None, 1.0, 0)
        return mu.unsqueeze(-1).unsqueeze(-1)


class _ZInitial(KernelModule, nn.Sequential):
    """z_initial = Conv1x1(L->512) + BN + ReLU (unet_resnet.py:150-154)."""

    def __init__(self, latent_dim):
        nn.Sequential.__init__(self, nn.Conv2d(latent_dim, 512, 1), nn.BatchNorm2d(512), nn.ReLU(inplace=True))

    def _build_unit(self, sd, dev):
        return ZInitialUnit(sd, "", dev, self.compute_dtype)

    def forward(self, z_spatial):
        self._require_eval()
        return from_act(self._unit()(to_act(z_spatial, self.compute_dtype)))


class _FinalConv(KernelModule, nn.Conv2d):
    """final_conv = nn.Conv2d(64, n_classes, 1) (unet_resnet.py:189) on the 1x1 head kernel."""

    def __init__(self, cin, n_classes):
        nn.Conv2d.__init__(self, cin, n_classes, kernel_size=1)

    def _build_unit(self, sd, dev):
        return FinalConvUnit(sd, "", dev)

    def forward(self, x):
        return self._unit()(to_act(x, self.compute_dtype))


class UNetResNet(KernelModule):
    """unet/unet_resnet.py:103-279. forward -> (logits [B,n_classes,H,W], mu [B,L], logvar [B,L]).
    Additive extension: `forward(x, eps=...)` injects the reparameterisation noise."""

    def __init__(self, n_channels, n_classes, backbone='resnet34', pretrained=True, latent_dim=32,
                 use_attention=True, use_skip=True, latent_injection='all'):
        super().__init__()
        self.n_channels = n_channels
        self.n_classes = n_classes
        self.latent_dim = latent_dim
        self.latent_injection = latent_injection
        if backbone != 'resnet34':
            raise ValueError(f"backbone '{backbone}' is not available: the reference hard-codes ResNet-18/34 decoder "
                             "widths (unet_resnet.py:150-186); this build ships resnet34")
        if pretrained:
            raise RuntimeError("pretrained=True downloads timm weights; there is no network here. "
                               "Construct with pretrained=False and load a state_dict.")
        self.encoder = ResNet34Features(in_chans=n_channels)
        encoder_channels = self.encoder.feature_info.channels()
        self.mu_head = _LatentHead(encoder_channels[-1], latent_dim)
        self.logvar_head = _LatentHead(encoder_channels[-1], latent_dim)
        self.z_initial = _ZInitial(latent_dim)
        use_latent_list, self.use_bottleneck = _latent_flags(latent_injection)
        self.use_skip = use_skip
        self.use_attention = use_attention and use_skip
        self.decoder_blocks = nn.ModuleList([
            DecoderBlock(512, encoder_channels[-2], 512, latent_dim, use_attention, use_skip, use_latent_list[0]),
            DecoderBlock(512, encoder_channels[-3], 256, latent_dim, use_attention, use_skip, use_latent_list[1]),
            DecoderBlock(256, encoder_channels[-4], 128, latent_dim, use_attention, use_skip, use_latent_list[2]),
            DecoderBlock(128, encoder_channels[0], 64, latent_dim, use_attention, use_skip, use_latent_list[3]),
        ])
        self.final_conv = _FinalConv(64, n_classes)

    # the whole-model engine (packed weights for every layer)
    def _build_unit(self, sd, dev):
        return ResNetVAEEngine(sd, dev, self.latent_injection, self.use_attention, self.use_skip, self.latent_dim,
                               self.n_channels, self.n_classes, self.compute_dtype)

    def engine(self) -> ResNetVAEEngine:
        return self._unit()

    def reparameterize(self, mu, logvar, eps=None):
        std = torch.exp(0.5 * logvar)
        if eps is None:
            eps = torch.randn_like(std)
        return mu + eps * std

    def forward(self, x, eps: Optional[torch.Tensor] = None):
        if self.training:
            # train mode (batch-statistics BatchNorm, autograd): torch.ops.vaeunet.* custom ops, any configuration
            from ..autograd_path import unetresnet_forward
            return unetresnet_forward(self, x, eps)
        return self.engine().forward(x.float().contiguous(), eps)

    def encode(self, x):
        self._require_eval()
        eng = self.engine()
        feats = eng.encode(x.float().contiguous())
        mu, lv, _ = eng.heads(feats[-1], None, 1.0, 0)
        return mu, lv

    def decode(self, z, input_size=None):
        """unet_resnet.py:250-279: skips come from the encoder run on one all-zero 512x512 image."""
        self._require_eval()
        eng = self.engine()
        feats = eng.encode(torch.zeros(1, self.n_channels, 512, 512, device=z.device))
        D = z.shape[0]
        if not self.use_bottleneck:
            f4 = feats[-1]
            feats = list(feats[:-1]) + [ops.Act(torch.zeros((1, f4.H, f4.W, f4.C), dtype=f4.t.dtype, device=z.device))]
        if D > 1 and not self.use_bottleneck:
            f4 = feats[-1]
            feats[-1] = ops.Act(f4.t.expand(1, -1, -1, -1).contiguous())
        lg = eng.decode(feats, z.float().contiguous())
        if input_size is not None:
            h, w = lg.shape[2:]
            lg = ops.resize_logits(lg.reshape(-1, h, w), int(input_size[0]), int(input_size[1]), ops.SIG_NONE)
            lg = lg.reshape(D, self.n_classes, int(input_size[0]), int(input_size[1]))
        return lg

    # additive fast path
    def mc_predict(self, images, temperature=1.0, num_samples=10, eps=None, return_samples=False,
                   outs=("mean", "std", "entropy", "mutual_info", "coeff_var")):
        """`outs`: which maps of calculate_uncertainty_metrics (visualize_vae.py:90-117) to produce."""
        self._require_eval()
        return self.engine().mc_predict(images.float().contiguous(), temperature, num_samples, eps, return_samples, outs)
