"""Inference engine for the ResNet-34-encoder VAE-UNet on B200.

Holds the reference's weights (same `state_dict` keys) re-packed for the sm_100a
kernels: K-major 16-bit conv filters with eval-mode BatchNorm folded into fp32 scale/bias
epilogue vectors. Activations stay NHWC 16-bit on the device between layers; `torch.cat` is
replaced by writing into channel slices of one buffer; the N Monte-Carlo decodes of a patch run as
ONE batched decoder pass (decoder batch index d = patch * Ns + sample) that shares the encoder
features — the "N-sample MC decode that reuses a single encoder pass" of BASELINE.json.

Reference behaviour mirrored here: unet/unet_resnet.py:31-101 (DecoderBlock), :196-240 (forward),
utils/vae_utils.py:28-76 (generate_predictions), visualize_vae.py:61-117 (predict_full_image,
calculate_uncertainty_metrics).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Union

import torch

from . import ops
from .ops import Act
from ._nvtx import rng

BN_EPS = 1e-5
RESNET34_STAGES = (3, 4, 6, 3)
# same switch the C side reads (vu_conv_halo.cu): with the tap-reuse kernel off there is no fused output conv either
_NO_HALO = os.environ.get("VU_NO_HALO", "0") not in ("", "0")


@dataclass
class ConstVec:
    """A spatially constant map, stored as one fp32 vector per image: v[D, C]. `hw` is the size the reference's map would
    have (needed only when no skip connection defines the output size of the next DecoderBlock: use_skip=False)."""
    v: torch.Tensor
    hw: Optional[tuple] = None


def _latent_flags(latent_injection):
    li = latent_injection
    if isinstance(li, list):
        return [i in li for i in range(4)], True
    known = ("all", "inject_no_bottleneck", "first", "last", "bottleneck", "none")
    if li not in known:
        li = "all"
    use = {"all": [True] * 4, "inject_no_bottleneck": [True] * 4, "first": [True, False, False, False],
           "last": [False, False, False, True], "bottleneck": [False] * 4, "none": [False] * 4}[li]
    return use, li not in ("none", "inject_no_bottleneck")


def fold_bn(sd: Dict[str, torch.Tensor], bn: str, conv_bias: Optional[torch.Tensor], dev) -> tuple:
    """Eval BatchNorm2d folded behind a conv: y = scale * conv(x) + bias."""
    g = sd[bn + "weight"].detach().to(dev, torch.float32)
    b = sd[bn + "bias"].detach().to(dev, torch.float32)
    m = sd[bn + "running_mean"].detach().to(dev, torch.float32)
    v = sd[bn + "running_var"].detach().to(dev, torch.float32)
    scale = g / torch.sqrt(v + BN_EPS)
    bias = b - m * scale
    if conv_bias is not None:
        bias = bias + conv_bias.detach().to(dev, torch.float32) * scale
    return scale.contiguous(), bias.contiguous()


class ConvUnit:
    """conv (k x k, stride, pad) + folded BN (+ReLU, +residual) on the tcgen05 implicit-GEMM kernel."""

    def __init__(self, sd, conv: str, bn: Optional[str], k: int, stride: int, pad: int, dev, dtype, cin_keep: Optional[int] = None):
        w = sd[conv + "weight"].detach().to(dev, torch.float32).contiguous()
        if cin_keep is not None:                      # filter bank over a slice of the input channels only
            lo, hi = (0, cin_keep) if isinstance(cin_keep, int) else cin_keep
            w = w[:, lo:hi].contiguous()
        cb = sd.get(conv + "bias")
        self.k, self.stride, self.pad = k, stride, pad
        self.cout, self.cin = w.shape[0], w.shape[1]
        self.w = ops.pack_conv_weight(w, dtype)
        if bn is not None:
            self.scale, self.bias = fold_bn(sd, bn, cb, dev)
        else:
            self.scale = None
            self.bias = cb.detach().to(dev, torch.float32).contiguous() if cb is not None else None

    def __call__(self, x: Act, relu: bool, out: Optional[Act] = None, residual: Optional[Act] = None, **kw) -> Act:
        return ops.conv2d(x, self.w, self.k, self.stride, self.pad, self.scale, self.bias, relu, out=out,
                          residual=residual, **kw)

    def can_fuse_output(self, x: Act) -> bool:
        """vu_conv_args.fuse_out: 3x3 s1 p1, Cout 64/128, map at least 96 wide."""
        return self.k == 3 and self.stride == 1 and self.cout in (64, 128) and x.W >= 96 and not _NO_HALO


class VecUnit:
    """1x1 conv + BN (+ReLU) applied to a spatially constant map = a tiny per-image mat-vec."""

    def __init__(self, sd, conv: str, bn: Optional[str], dev):
        self.w = sd[conv + "weight"].detach().to(dev, torch.float32).reshape(sd[conv + "weight"].shape[0], -1).contiguous()
        cb = sd.get(conv + "bias")
        if bn is not None:
            self.scale, self.shift = fold_bn(sd, bn, cb, dev)
        else:
            self.scale = None
            self.shift = cb.detach().to(dev, torch.float32).contiguous() if cb is not None else None

    def __call__(self, z: torch.Tensor, relu: bool) -> torch.Tensor:
        return ops.latent_project(z.contiguous(), self.w, self.scale, self.shift, relu)


class GateUnit:
    """AttentionGate (unet/unet_resnet.py:6-29): two tensor-core 1x1 GEMMs + one fused gate kernel."""

    def __init__(self, sd, p: str, dev, dtype):
        self.wg = ConvUnit(sd, p + "W_g.0.", p + "W_g.1.", 1, 1, 0, dev, dtype)
        self.wx = ConvUnit(sd, p + "W_x.0.", p + "W_x.1.", 1, 1, 0, dev, dtype)
        self.wg_vec = VecUnit(sd, p + "W_g.0.", p + "W_g.1.", dev)
        pw = sd[p + "psi.0.weight"].detach().to(dev, torch.float32).reshape(-1).contiguous()
        sc, bi = fold_bn(sd, p + "psi.1.", sd.get(p + "psi.0.bias"), dev)
        self.psi_w = pw
        self.psi_scale = float(sc.item())
        self.psi_bias = float(bi.item())
        self.fint = pw.numel()

    def __call__(self, g: Union[Act, ConstVec], skip: Act, out: Act, psi_out=None) -> Act:
        x1 = self.wx(skip, relu=False)
        if isinstance(g, ConstVec):
            g1 = ops.new_act(out.N, out.H, out.W, self.fint, out.dt, out.t.device)
            ops.broadcast_vec(self.wg_vec(g.v, relu=False), g1)
        else:
            g1 = self.wg(g, relu=False)
        return ops.attention_gate(g1, x1, skip, self.psi_w, self.psi_scale, self.psi_bias, out, psi_out)


class EncoderUnit:
    """ResNet-34 features_only (timm call site unet/unet_resnet.py:131-137): [act1, layer1..layer4]."""

    def __init__(self, sd, prefix: str, dev, dtype):
        p = prefix
        self.dtype = dtype
        self.stem_w = sd[p + "conv1.weight"].detach().to(dev, torch.float32).contiguous()
        self.stem_scale, self.stem_bias = fold_bn(sd, p + "bn1.", None, dev)
        self.blocks = []
        for li, nb in enumerate(RESNET34_STAGES, start=1):
            for b in range(nb):
                q = f"{p}layer{li}.{b}."
                stride = 2 if (li > 1 and b == 0) else 1
                c1 = ConvUnit(sd, q + "conv1.", q + "bn1.", 3, stride, 1, dev, dtype)
                c2 = ConvUnit(sd, q + "conv2.", q + "bn2.", 3, 1, 1, dev, dtype)
                ds = None
                if (q + "downsample.0.weight") in sd:
                    ds = ConvUnit(sd, q + "downsample.0.", q + "downsample.1.", 1, stride, 0, dev, dtype)
                self.blocks.append((li, c1, c2, ds))

    def __call__(self, images: torch.Tensor) -> List[Act]:
        """images NCHW fp32 on device -> [f0..f4] NHWC 16-bit."""
        # fp32 direct stem: rounding the input image to 16 bits (the tensor-core im2col route the training step uses) costs
        # 3x in output parity on the seeded fixture for a 3 % gain, so inference keeps full-precision pixels
        x = ops.stem_conv7x7s2(images, self.stem_w, self.stem_scale, self.stem_bias, self.dtype)
        feats = [x]
        x = ops.maxpool(x, 3, 2, 1)
        cur = 1
        for (li, c1, c2, ds) in self.blocks:
            if li != cur:
                feats.append(x)
                cur = li
            idt = ds(x, relu=False) if ds is not None else x
            y = c1(x, relu=True)
            x = c2(y, relu=True, residual=idt)
        feats.append(x)
        return feats


class HeadUnit:
    """mu_head / logvar_head: conv1x1 + GAP (unet/unet_resnet.py:140-147)."""

    def __init__(self, sd, prefix: str, dev):
        w = sd[prefix + "0.weight"].detach().to(dev, torch.float32)
        self.w = w.reshape(w.shape[0], -1).contiguous()
        self.b = sd[prefix + "0.bias"].detach().to(dev, torch.float32).contiguous()


class ZInitialUnit:
    """z_initial: 1x1 conv 32->512 + BN + ReLU (unet/unet_resnet.py:150-154)."""

    def __init__(self, sd, prefix: str, dev, dtype):
        self.sd, self.prefix, self.dev, self.dtype = sd, prefix, dev, dtype
        self.vec = VecUnit(sd, prefix + "0.", prefix + "1.", dev)
        self.conv = None

    def __call__(self, z: Union[ConstVec, Act]) -> Union[ConstVec, Act]:
        if isinstance(z, ConstVec):
            return ConstVec(self.vec(z.v, relu=True), z.hw)
        if self.conv is None:
            self.conv = ConvUnit(self.sd, self.prefix + "0.", self.prefix + "1.", 1, 1, 0, self.dev, self.dtype)
        return self.conv(z, relu=True)


class DecoderBlockUnit:
    """DecoderBlock.forward (unet/unet_resnet.py:71-101). x / z may be constant maps.
    skip.N may be smaller than the decoder batch D (D = skip.N * rep): MC samples share it."""

    def __init__(self, sd, prefix: str, dev, dtype, latent_dim: int, use_attention: bool, use_skip: bool,
                 use_latent: bool):
        p = prefix
        self.sd, self.prefix, self.dev, self.dtype, self.L = sd, p, dev, dtype, latent_dim
        self.use_skip = use_skip
        self.conv1 = ConvUnit(sd, p + "conv1.0.", p + "conv1.1.", 3, 1, 1, dev, dtype)
        self.conv2 = ConvUnit(sd, p + "conv2.0.", p + "conv2.1.", 3, 1, 1, dev, dtype)
        self.gate = GateUnit(sd, p + "attention.", dev, dtype) if (use_attention and use_skip) else None
        self.zproj = VecUnit(sd, p + "z_proj.0.", p + "z_proj.1.", dev) if use_latent else None
        self.zproj_conv = None
        # Constant-latent fast path: z_proj(z) is spatially constant per decode, so its 3x3 conv contribution is a per-decode
        # bias that only depends on the border class of the output pixel (zero padding). conv1 then runs over [x | skip_att]
        # only (224 -> 192, 352 -> 320, 672 -> 640, 800 -> 768 channels: whole 64-channel K blocks) and the concat buffer
        # loses its z slice; the z term is added in fp32 in the epilogue (ops.border_bias / conv2d(pos_bias=)).
        self.conv1_xs = None
        if use_latent:
            self.w1_full = sd[p + "conv1.0.weight"].detach().to(dev, torch.float32).contiguous()
            self.cin_xs = self.w1_full.shape[1] - latent_dim
            if self.cin_xs % 8 == 0 and self.cin_xs > 0:
                self.conv1_xs = ConvUnit(sd, p + "conv1.0.", p + "conv1.1.", 3, 1, 1, dev, dtype, cin_keep=self.cin_xs)
                self.wsum_z = ops.border_wsum(self.w1_full, self.cin_xs, latent_dim)
        self.conv1_s = None          # built on first use: constant x (block 0 fed by z_initial(z)) leaves conv1 as well

    def __call__(self, x: Union[Act, ConstVec], skip: Optional[Act], z: Union[ConstVec, Act, None], psi_out=None,
                 final: Optional["FinalConvUnit"] = None):
        """-> Act, or (with `final`, n_classes == 1) the fp32 logits [D,1,H,W] of final(conv2(...)) from the fused epilogue."""
        L = self.L
        if isinstance(x, ConstVec):
            D, cx = x.v.shape
        else:
            D, cx = x.N, x.C
        if skip is not None:
            H, W = skip.H, skip.W
        elif isinstance(x, Act):
            H, W = x.H * 2, x.W * 2                       # unet_resnet.py:75-76
        else:
            # use_skip=False with a bottleneck latent: x = z_initial(z_spatial) is a constant map of the size of the smallest
            # encoder feature (unet_resnet.py:221-229), doubled by this block like any other input
            assert x.hw is not None, "constant x without a skip needs its map size (ConstVec.hw)"
            H, W = x.hw[0] * 2, x.hw[1] * 2
        has_skip = skip is not None and self.use_skip
        cs = skip.C if has_skip else 0
        cz = L if self.zproj is not None else 0
        fused = has_skip and self.gate is not None and psi_out is None and (cz == 0 or isinstance(z, ConstVec))
        split_z = fused and cz > 0 and self.conv1_xs is not None and cx + cs == self.cin_xs and H >= 2 and W >= 2
        if split_z and isinstance(x, ConstVec) and cs % 8 == 0:
            # block 0 of the default model: x = z_initial(z) is constant too, so conv1 only sees the gated skip (800 -> 256
            # channels); x and z_proj(z) enter through the border-class bias
            if self.conv1_s is None:
                self.conv1_s = ConvUnit(self.sd, self.prefix + "conv1.0.", self.prefix + "conv1.1.", 3, 1, 1, self.dev, self.dtype,
                                        cin_keep=(cx, cx + cs))
                self.wsum_x = ops.border_wsum(self.w1_full, 0, cx)
            cat = ops.new_act(D, H, W, cs, self.dtype, self.dev)
            self.gate(x, skip, cat)
            pb = ops.border_bias_apply(self.wsum_x, x.v)
            ops.border_bias_apply(self.wsum_z, self.zproj(z.v, relu=True), out=pb)
            y = self.conv1_s(cat, relu=True, pos_bias=pb)
            return self._tail(y, final)
        cat = ops.new_act(D, H, W, cx + cs + (0 if split_z else cz), self.dtype, self.dev)
        if fused:
            # fused producer: W_g runs at the LOW resolution of x, everything else is one pass over `cat`
            gt = self.gate
            if isinstance(x, ConstVec):
                x_low = ops.broadcast_vec(x.v, ops.new_act(D, 1, 1, cx, self.dtype, self.dev))
                g1_low = ops.broadcast_vec(gt.wg_vec(x.v, relu=False), ops.new_act(D, 1, 1, gt.fint, self.dtype, self.dev))
            else:
                x_low = x
                g1_low = gt.wg(x, relu=False)
            x1 = gt.wx(skip, relu=False)
            zv = self.zproj(z.v, relu=True) if cz else None
            if split_z:
                ops.decoder_concat(x_low, g1_low, x1, skip, gt.psi_w, gt.psi_scale, gt.psi_bias, None, cat)
                return self._convs(cat, final, pos_bias=ops.border_bias_apply(self.wsum_z, zv))
            ops.decoder_concat(x_low, g1_low, x1, skip, gt.psi_w, gt.psi_scale, gt.psi_bias, zv, cat)
            return self._convs(cat, final)
        xs = cat.slice(0, cx)
        if isinstance(x, ConstVec):
            ops.broadcast_vec(x.v, xs)
            g = x
        else:
            ops.bilinear_ac(x, H, W, out=xs)
            g = xs
        if has_skip:
            ss = cat.slice(cx, cs)
            if self.gate is not None:
                self.gate(g, skip, ss, psi_out)
            else:
                rep = D // skip.N
                src = skip if rep == 1 else Act(skip.t[..., skip.coff:skip.coff + skip.C].repeat_interleave(rep, 0).contiguous())
                ops.bilinear_ac(src, H, W, out=ss)
        if cz:
            zs = cat.slice(cx + cs, cz)
            if isinstance(z, ConstVec):
                ops.broadcast_vec(self.zproj(z.v, relu=True), zs)
            else:
                zu = ops.bilinear_ac(z, H, W)
                if self.zproj_conv is None:
                    self.zproj_conv = ConvUnit(self.sd, self.prefix + "z_proj.0.", self.prefix + "z_proj.1.", 1, 1, 0,
                                               self.dev, self.dtype)
                self.zproj_conv(zu, relu=True, out=zs)
        return self._convs(cat, final)

    def _convs(self, cat: Act, final, pos_bias=None):
        y = self.conv1(cat, relu=True) if pos_bias is None else self.conv1_xs(cat, relu=True, pos_bias=pos_bias)
        return self._tail(y, final)

    def _tail(self, y: Act, final):
        if final is not None and final.n_classes == 1 and self.conv2.can_fuse_output(y):
            logits = torch.empty((y.N, 1, y.H, y.W), dtype=torch.float32, device=self.dev)
            self.conv2(y, relu=True, fuse_w=final.w[0], fuse_bias=final.b[0], fuse_out=logits, store=False)
            return logits
        x = self.conv2(y, relu=True)
        return x if final is None else final(x)


class FinalConvUnit:
    """final_conv / OutConv: 1x1 conv C->n_classes + bias, fp32 logits out."""

    def __init__(self, sd, prefix: str, dev):
        fw = sd[prefix + "weight"].detach().to(dev, torch.float32)
        self.n_classes = fw.shape[0]
        self.w = [fw[c].reshape(-1).contiguous() for c in range(self.n_classes)]
        self.b = [float(v) for v in sd[prefix + "bias"].detach().float().cpu()]

    def __call__(self, x: Act) -> torch.Tensor:
        outs = [ops.final_conv1(x, self.w[c], self.b[c]) for c in range(self.n_classes)]
        return outs[0].unsqueeze(1) if len(outs) == 1 else torch.stack(outs, 1)


class ResNetVAEEngine:
    def __init__(self, sd: Dict[str, torch.Tensor], device, latent_injection="all", use_attention=True,
                 use_skip=True, latent_dim=32, n_channels=3, n_classes=1, dtype=ops.DEFAULT_DTYPE):
        ops.require_device()
        self.dev = torch.device(device)
        self.dtype = dtype
        self.latent_dim = latent_dim
        self.n_channels, self.n_classes = n_channels, n_classes
        self.latent_injection = latent_injection
        self.use_latent, self.use_bottleneck = _latent_flags(latent_injection)
        self.use_skip = use_skip
        self.use_attention = use_attention and use_skip
        dev = self.dev
        self.encoder = EncoderUnit(sd, "encoder.", dev, dtype)
        self.mu_head = HeadUnit(sd, "mu_head.", dev)
        self.lv_head = HeadUnit(sd, "logvar_head.", dev)
        self.z_initial = ZInitialUnit(sd, "z_initial.", dev, dtype)
        self.dec = [DecoderBlockUnit(sd, f"decoder_blocks.{i}.", dev, dtype, latent_dim, self.use_attention,
                                     use_skip, self.use_latent[i]) for i in range(4)]
        self.final = FinalConvUnit(sd, "final_conv.", dev)

    def encode(self, images: torch.Tensor) -> List[Act]:
        with rng("vaeunet.encode"):
            return self.encoder(images)

    def heads(self, f4: Act, eps: Optional[torch.Tensor], temperature: float, Ns: int):
        with rng("vaeunet.latent_heads"):
            return self._heads(f4, eps, temperature, Ns)

    def _heads(self, f4: Act, eps: Optional[torch.Tensor], temperature: float, Ns: int):
        return ops.latent_head(f4, self.mu_head.w, self.mu_head.b, self.lv_head.w, self.lv_head.b, eps, temperature, Ns)

    def decode(self, feats: Sequence[Act], z_dec: torch.Tensor, honour_use_skip: bool = True) -> torch.Tensor:
        """z_dec fp32 [D, L] with D = B * rep (sample-minor) -> logits fp32 [D, n_classes, H/2, W/2]."""
        with rng("vaeunet.decode"):
            return self._decode(feats, z_dec, honour_use_skip)

    def _decode(self, feats: Sequence[Act], z_dec: torch.Tensor, honour_use_skip: bool = True) -> torch.Tensor:
        z = ConstVec(z_dec.contiguous(), (feats[-1].H, feats[-1].W))
        D = z_dec.shape[0]
        if self.use_bottleneck:
            x: Union[Act, ConstVec] = self.z_initial(z)
        else:
            f4 = feats[-1]
            rep = D // f4.N
            x = f4 if rep == 1 else Act(f4.t.repeat_interleave(rep, 0).contiguous(), f4.coff, f4.C)
        for i in range(4):
            want = self.use_skip if honour_use_skip else True
            skip = feats[-(i + 2)] if want else None
            x = self.dec[i](x, skip, z, final=self.final if i == 3 else None)
        return x

    # ------------------------------------------------------------------ end-to-end entry points
    def forward(self, images: torch.Tensor, eps: Optional[torch.Tensor] = None):
        """UNetResNet.forward in eval mode (unet/unet_resnet.py:196-240). eps [B,L]."""
        B, _, H, W = images.shape
        feats = self.encode(images)
        sample = self.latent_injection not in ("none", "inject_no_bottleneck")
        e = eps.reshape(1, B, self.latent_dim).contiguous() if (sample and eps is not None) else None
        if sample and e is None:
            e = torch.randn(1, B, self.latent_dim, device=self.dev)
        mu, lv, z = self.heads(feats[-1], e if sample else None, 1.0, 1)
        lg = self.decode(feats, z[0])
        out = ops.resize_logits(lg.reshape(-1, lg.shape[2], lg.shape[3]), H, W, ops.SIG_NONE)
        return out.reshape(B, self.n_classes, H, W), mu, lv

    def mc_logits_honour_skip(self, images, temperature, Ns, eps, sample=True):
        """generate_predictions flavour (utils/vae_utils.py:63-65 honours model.use_skip)."""
        return self.mc_logits(images, temperature, Ns, eps, sample, honour_use_skip=True)

    def mc_logits(self, images: torch.Tensor, temperature: float, Ns: int, eps: Optional[torch.Tensor],
                  sample: bool = True, honour_use_skip: bool = False):
        """One encoder pass, Ns batched decodes. -> logits [B, Ns, h, w] (H/2), mu, logvar."""
        B = images.shape[0]
        feats = self.encode(images)
        if sample and eps is None:
            eps = torch.randn(Ns, B, self.latent_dim, device=self.dev)
        mu, lv, z = self.heads(feats[-1], eps if sample else None, temperature, Ns)
        z_dec = z.permute(1, 0, 2).reshape(B * Ns, self.latent_dim)   # sample-minor
        lg = self.decode(feats, z_dec, honour_use_skip=honour_use_skip)
        return lg.reshape(B, Ns, lg.shape[2], lg.shape[3]), mu, lv

    def mc_predict(self, images: torch.Tensor, temperature: float = 1.0, num_samples: int = 10,
                   eps: Optional[torch.Tensor] = None, return_samples: bool = False,
                   outs=("mean", "std", "entropy", "mutual_info", "coeff_var")) -> Dict[str, torch.Tensor]:
        """BASELINE config 2: per patch mean / std / entropy / mutual_info / coeff_var maps over
        `num_samples` latent draws (predict_full_image order: upsample, then sigmoid)."""
        assert self.n_classes == 1
        H, W = images.shape[2:]
        sample = self.latent_injection != "none"
        lg, mu, lv = self.mc_logits(images, temperature, num_samples, eps, sample)
        with rng("vaeunet.uncertainty_reduce"):
            res = ops.uncertainty_reduce(lg.contiguous(), H, W, want_probs=return_samples, outs=tuple(outs))
        res["mu"], res["logvar"] = mu, lv
        return res


class GraphedMC:
    """`mc_predict` for one fixed (batch, size, N, T) captured in a CUDA graph: the ~70 launches of a step replay without
    per-launch host work or inter-kernel gaps. Inputs are copied into static buffers, outputs are static tensors that the
    next replay overwrites (copy them out if they must survive). Tensor maps are encoded once, at capture."""

    def __init__(self, eng: "ResNetVAEEngine", batch: int, height: int, width: int, num_samples: int = 10,
                 temperature: float = 1.0, warmup: int = 2, outs=("mean", "std", "entropy", "mutual_info", "coeff_var")):
        self.eng = eng
        self.outs = tuple(outs)
        dev = eng.dev
        self.x = torch.zeros((batch, eng.n_channels, height, width), dtype=torch.float32, device=dev)
        self.eps = torch.zeros((num_samples, batch, eng.latent_dim), dtype=torch.float32, device=dev)
        self.T, self.N = temperature, num_samples
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(warmup):                      # first-call initialisation (function attributes, caches) stays outside
                eng.mc_predict(self.x, temperature, num_samples, self.eps, outs=self.outs)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        l0 = ops.LAUNCHES
        with torch.cuda.graph(self.graph), torch.no_grad():
            self.out = eng.mc_predict(self.x, temperature, num_samples, self.eps, outs=self.outs)
        self.launches = ops.LAUNCHES - l0           # kernels of this repo inside one replay

    def __call__(self, images: torch.Tensor, eps: torch.Tensor) -> Dict[str, torch.Tensor]:
        self.x.copy_(images, non_blocking=True)
        self.eps.copy_(eps, non_blocking=True)
        self.graph.replay()
        ops._count(self.launches)
        return self.out
