"""CPU oracle for the VAE-UNet hot path (TEST INFRASTRUCTURE, NOT PRODUCT CODE).

A plain PyTorch fp32 *functional* restatement of the reference's algorithm for the
hot path named by BASELINE.json: the ResNet-34-encoder VAE-UNet (`UNetResNet`), the
basic attention U-Net (`UNet`), the Monte-Carlo latent sampling helpers, the per-pixel
uncertainty reduction, the tiled/blended full-image predictor, the losses and the scalar
uncertainty metrics.  Every function cites the reference file:line it follows
(paths relative to the upstream repo tmuird/VAEUNET).

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference`
legs may import this module.  The product package `vaeunet_b200` never does.

Parity status: PINNED.  `tests/golden/*.pt` were produced by importing the reference's own
modules in the build container (`tests/golden/make_golden.py`; two import shims: a fake
`timm` whose `create_model` builds a torchvision ResNet-34 feature net — timm ~=1.0.13 is an
un-vendored dependency, `requirements.txt:2`, call site `unet/unet_resnet.py:131-137` — and an
empty `matplotlib`).  `tests/test_oracle_golden.py` checks this restatement against those
vectors.  The encoder arithmetic itself (timm resnet34 features_only) is the published
ResNet-34 architecture: 7x7 s2 p3 stem conv + BN + ReLU, 3x3 s2 p1 max-pool, BasicBlock
stages [3,4,6,3] with 1x1-s2-conv+BN shortcuts, BN eps 1e-5; features are
[act1, layer1, layer2, layer3, layer4].

All weights come from a `state_dict` with the reference's own key layout.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
SD = Dict[str, Tensor]

BN_EPS = 1e-5
BN_MOMENTUM = 0.1
RESNET34_STAGES = (3, 4, 6, 3)


# --------------------------------------------------------------------------------------
# configuration mirrors of the reference constructors
# --------------------------------------------------------------------------------------
@dataclass
class ResNetCfg:
    """Mirror of UNetResNet.__init__ flags (unet/unet_resnet.py:104-186)."""
    n_channels: int = 3
    n_classes: int = 1
    latent_dim: int = 32
    use_attention: bool = True
    use_skip: bool = True
    latent_injection: object = "all"

    @property
    def use_latent_list(self) -> List[bool]:
        li = self.latent_injection
        # unet/unet_resnet.py:157-172
        if isinstance(li, list):
            return [i in li for i in range(4)]
        if li in ("all", "inject_no_bottleneck"):
            return [True] * 4
        if li == "first":
            return [True, False, False, False]
        if li == "last":
            return [False, False, False, True]
        if li in ("bottleneck", "none"):
            return [False] * 4
        return [True] * 4

    @property
    def use_bottleneck(self) -> bool:
        # unet/unet_resnet.py:175 (note: evaluated on the *corrected* mode, so an unknown
        # string behaves like 'all')
        li = self.latent_injection
        if not isinstance(li, list) and li not in (
                "all", "inject_no_bottleneck", "first", "last", "bottleneck", "none"):
            li = "all"
        return li not in ("none", "inject_no_bottleneck")

    @property
    def attention_on(self) -> bool:
        return self.use_attention and self.use_skip  # unet/unet_resnet.py:179


@dataclass
class BNState:
    """Train-mode switch + collector of running-stat updates (nn.BatchNorm2d semantics)."""
    training: bool = False
    updates: Dict[str, Tensor] = field(default_factory=dict)


# --------------------------------------------------------------------------------------
# primitive layers
# --------------------------------------------------------------------------------------
def _bn(sd: SD, p: str, x: Tensor, st: Optional[BNState]) -> Tensor:
    """nn.BatchNorm2d(eps=1e-5, momentum=0.1). Eval: running stats. Train: batch stats and
    running-stat update recorded in st.updates (does not mutate sd)."""
    if st is not None and st.training:
        rm = sd[p + "running_mean"].clone()
        rv = sd[p + "running_var"].clone()
        y = F.batch_norm(x, rm, rv, sd[p + "weight"], sd[p + "bias"], True, BN_MOMENTUM, BN_EPS)
        st.updates[p + "running_mean"] = rm
        st.updates[p + "running_var"] = rv
        return y
    return F.batch_norm(x, sd[p + "running_mean"], sd[p + "running_var"],
                        sd[p + "weight"], sd[p + "bias"], False, BN_MOMENTUM, BN_EPS)


# --------------------------------------------------------------------------------------
# optional emulation of 16-bit activation / weight STORAGE (tests only)
# --------------------------------------------------------------------------------------
# The B200 kernels keep activations and tensor-core filter banks as bf16 or fp16 and do all
# arithmetic in fp32. With STORAGE_DTYPE set, `_q` rounds at exactly the points where the kernels
# write a 16-bit buffer, so a GPU result can be checked against "the reference algorithm with the
# same storage rounding" (kernel correctness) separately from "fp32 reference" (rounding budget).
# The golden-vector tests always run with STORAGE_DTYPE = None.
STORAGE_DTYPE: Optional[torch.dtype] = None


class storage_rounding:
    """with storage_rounding(torch.bfloat16): ... — round at the kernels' 16-bit storage points."""

    def __init__(self, dtype: Optional[torch.dtype]):
        self.dtype = dtype

    def __enter__(self):
        global STORAGE_DTYPE
        self.prev = STORAGE_DTYPE
        STORAGE_DTYPE = self.dtype
        return self

    def __exit__(self, *exc):
        global STORAGE_DTYPE
        STORAGE_DTYPE = self.prev
        return False


def _q(x: Tensor) -> Tensor:
    return x if STORAGE_DTYPE is None else x.to(STORAGE_DTYPE).to(torch.float32)


def _is_const_map(x: Tensor) -> bool:
    return bool((x.amax(dim=(2, 3)) == x.amin(dim=(2, 3))).all())


def _conv(sd: SD, p: str, x: Tensor, stride: int = 1, padding: int = 0, fp32: bool = False) -> Tensor:
    """fp32=True marks the convs the kernels run with fp32 weights (stem, heads, output conv and the
    1x1 convs on spatially constant latent maps); all others use 16-bit filter banks."""
    w = sd[p + "weight"]
    if not fp32:
        w = _q(w)
    return F.conv2d(x, w, sd.get(p + "bias"), stride=stride, padding=padding)


def _up_ac(x: Tensor, size) -> Tensor:
    """F.interpolate(mode='bilinear', align_corners=True) as used throughout the reference
    (unet/unet_resnet.py:79,93,221,238)."""
    return F.interpolate(x, size=tuple(size), mode="bilinear", align_corners=True)


# --------------------------------------------------------------------------------------
# encoder: ResNet-34 features_only (timm call site unet/unet_resnet.py:131-137,201)
# --------------------------------------------------------------------------------------
def resnet34_features(sd: SD, x: Tensor, prefix: str = "encoder.",
                      st: Optional[BNState] = None) -> List[Tensor]:
    p = prefix
    x = _q(F.relu(_bn(sd, p + "bn1.", _conv(sd, p + "conv1.", x, 2, 3, fp32=True), st)))
    feats = [x]
    x = F.max_pool2d(x, 3, 2, 1)
    for li, nblk in enumerate(RESNET34_STAGES, start=1):
        for b in range(nblk):
            q = f"{p}layer{li}.{b}."
            stride = 2 if (li > 1 and b == 0) else 1
            idt = x
            y = _q(F.relu(_bn(sd, q + "bn1.", _conv(sd, q + "conv1.", x, stride, 1), st)))
            y = _bn(sd, q + "bn2.", _conv(sd, q + "conv2.", y, 1, 1), st)
            if (q + "downsample.0.weight") in sd:
                idt = _q(_bn(sd, q + "downsample.1.", _conv(sd, q + "downsample.0.", x, stride, 0), st))
            x = _q(F.relu(y + idt))
        feats.append(x)
    return feats


# --------------------------------------------------------------------------------------
# attention gate / decoder block / heads
# --------------------------------------------------------------------------------------
def attention_gate(sd: SD, p: str, g: Tensor, x: Tensor, st: Optional[BNState] = None,
                   return_psi: bool = False, g_fp32: bool = False):
    """x * sigmoid(BN(psi(ReLU(BN(W_g g) + BN(W_x x)))))
    unet/unet_parts.py:7-30 == unet/unet_resnet.py:6-29."""
    g1 = _q(_bn(sd, p + "W_g.1.", _conv(sd, p + "W_g.0.", g, fp32=g_fp32), st))
    x1 = _q(_bn(sd, p + "W_x.1.", _conv(sd, p + "W_x.0.", x), st))
    a = F.relu(g1 + x1)
    psi = torch.sigmoid(_bn(sd, p + "psi.1.", _conv(sd, p + "psi.0.", a, fp32=True), st))
    return (_q(x * psi), psi) if return_psi else _q(x * psi)


def decoder_block(sd: SD, p: str, x: Tensor, skip: Optional[Tensor], z: Tensor,
                  use_attention: bool, use_skip: bool, use_latent: bool,
                  st: Optional[BNState] = None) -> Tensor:
    """DecoderBlock.forward, unet/unet_resnet.py:71-101. Concat order [x, skip_att, z_proj]."""
    if skip is not None:
        size = skip.shape[2:]
    else:
        size = (x.shape[2] * 2, x.shape[3] * 2)
    # storage emulation: a spatially constant x (z_initial output) stays an fp32 vector in the
    # kernels; it is rounded only where it is written into the concat buffer
    x_const = STORAGE_DTYPE is not None and _is_const_map(x)
    x = _up_ac(x, size)
    g = x if x_const else _q(x)
    comps = [_q(x)]
    if skip is not None and use_skip:
        if use_attention:
            skip = attention_gate(sd, p + "attention.", g, skip, st, g_fp32=x_const)
        comps.append(skip)
    if use_latent:
        zp = _up_ac(z, size)
        z_const = STORAGE_DTYPE is not None and _is_const_map(zp)
        zp = zp if (z_const or STORAGE_DTYPE is None) else _q(zp)
        zp = _q(F.relu(_bn(sd, p + "z_proj.1.", _conv(sd, p + "z_proj.0.", zp, fp32=z_const), st)))
        comps.append(zp)
    x = torch.cat(comps, dim=1)
    x = _q(F.relu(_bn(sd, p + "conv1.1.", _conv(sd, p + "conv1.0.", x, 1, 1), st)))
    x = _q(F.relu(_bn(sd, p + "conv2.1.", _conv(sd, p + "conv2.0.", x, 1, 1), st)))
    return x


def latent_heads(sd: SD, x_enc: Tensor) -> Tuple[Tensor, Tensor]:
    """mu/logvar = GAP(conv1x1(x_enc)); unet/unet_resnet.py:140-147,205-206."""
    mu = F.adaptive_avg_pool2d(_conv(sd, "mu_head.0.", x_enc, fp32=True), 1).squeeze(-1).squeeze(-1)
    lv = F.adaptive_avg_pool2d(_conv(sd, "logvar_head.0.", x_enc, fp32=True), 1).squeeze(-1).squeeze(-1)
    return mu, lv


def sample_from_latent(mu: Tensor, logvar: Tensor, temperature: float, eps: Tensor) -> Tensor:
    """utils/vae_utils.py:5-10 with the noise injected: z = mu + eps * exp(.5 logvar) * T."""
    std = torch.exp(0.5 * logvar) * temperature
    return mu + eps * std


def decode_from_features(sd: SD, cfg: ResNetCfg, features: Sequence[Tensor], z: Tensor,
                         honour_use_skip: bool = True, st: Optional[BNState] = None) -> Tensor:
    """z [B,L] (or [B,L,1,1]) + encoder features -> logits at H/2 (before the final upsample).
    unet/unet_resnet.py:217-237; utils/vae_utils.py:53-67; visualize_vae.py:70-79."""
    if z.dim() == 2:
        z = z.unsqueeze(-1).unsqueeze(-1)
    z_sp = _up_ac(z, features[-1].shape[2:])
    if cfg.use_bottleneck:
        x = F.relu(_bn(sd, "z_initial.1.", _conv(sd, "z_initial.0.", z_sp, fp32=True), st))
    else:
        x = features[-1]
    ul = cfg.use_latent_list
    for i in range(4):
        want_skip = cfg.use_skip if honour_use_skip else True
        skip = features[-(i + 2)] if (i < len(features) - 1 and want_skip) else None
        x = decoder_block(sd, f"decoder_blocks.{i}.", x, skip, z_sp,
                          cfg.attention_on, cfg.use_skip, ul[i], st)
    return _conv(sd, "final_conv.", x, fp32=True)


# --------------------------------------------------------------------------------------
# UNetResNet entry points
# --------------------------------------------------------------------------------------
def unetresnet_forward(sd: SD, cfg: ResNetCfg, x: Tensor, eps: Optional[Tensor] = None,
                       st: Optional[BNState] = None) -> Tuple[Tensor, Tensor, Tensor]:
    """UNetResNet.forward, unet/unet_resnet.py:196-240. `eps` [B,L] replaces randn_like."""
    feats = resnet34_features(sd, x, "encoder.", st)
    mu, logvar = latent_heads(sd, feats[-1])
    if cfg.latent_injection not in ("none", "inject_no_bottleneck"):  # :210
        assert eps is not None, "eps must be injected for parity"
        z = mu + eps * torch.exp(0.5 * logvar)  # :191-194
    else:
        z = mu
    out = decode_from_features(sd, cfg, feats, z, True, st)
    out = _up_ac(out, x.shape[2:])
    return out, mu, logvar


def unetresnet_encode(sd: SD, x: Tensor) -> Tuple[Tensor, Tensor]:
    """UNetResNet.encode (unet/unet_resnet.py:242-248) == vae_utils.encode_images (:13-25)."""
    return latent_heads(sd, resnet34_features(sd, x)[-1])


def unetresnet_decode(sd: SD, cfg: ResNetCfg, z: Tensor, input_size=None) -> Tensor:
    """UNetResNet.decode, unet/unet_resnet.py:250-279: skips come from an all-zero 512x512
    image of batch 1 (:253-254); without bottleneck x starts from zeros (:267)."""
    feats = resnet34_features(sd, torch.zeros(1, cfg.n_channels, 512, 512))
    zz = z.unsqueeze(-1).unsqueeze(-1)
    z_sp = _up_ac(zz, feats[-1].shape[2:])
    if cfg.use_bottleneck:
        x = F.relu(_bn(sd, "z_initial.1.", _conv(sd, "z_initial.0.", z_sp, fp32=True), None))
    else:
        x = torch.zeros(z.size(0), 512, *feats[-1].shape[2:])
    ul = cfg.use_latent_list
    for i in range(4):
        skip = feats[-(i + 2)] if cfg.use_skip else None
        x = decoder_block(sd, f"decoder_blocks.{i}.", x, skip, z_sp,
                          cfg.attention_on, cfg.use_skip, ul[i], None)
    out = _conv(sd, "final_conv.", x, fp32=True)
    if input_size is not None:
        out = _up_ac(out, input_size)
    return out


def generate_predictions(sd: SD, cfg: ResNetCfg, images: Tensor, temperature: float,
                         num_samples: int, eps: Optional[Tensor]) -> Tensor:
    """utils/vae_utils.py:28-76: one encoder pass, N decodes, mean of LOGITS at H/2.
    eps [N,B,L]; the reference scales eps by T (:49) then by std (:50)."""
    feats = resnet34_features(sd, images)
    mu, logvar = latent_heads(sd, feats[-1])
    should_sample = cfg.latent_injection != "none"  # :43
    preds = []
    for n in range(num_samples):
        if should_sample:
            std = torch.exp(0.5 * logvar)
            z = mu + (eps[n] * temperature) * std
        else:
            z = mu
        preds.append(decode_from_features(sd, cfg, feats, z, True))
    return torch.stack(preds).mean(0) if len(preds) > 1 else preds[0]


def predict_full_image(sd: SD, cfg: ResNetCfg, img: Tensor, z: Tensor) -> Tensor:
    """visualize_vae.py:61-87: encoder + decode one z, upsample, THEN sigmoid.
    Ignores model.use_skip when picking skips (:76)."""
    feats = resnet34_features(sd, img)
    out = decode_from_features(sd, cfg, feats, z, honour_use_skip=False)
    return torch.sigmoid(_up_ac(out, img.shape[2:]))


def patch_grid(H: int, W: int, patch_size: int, overlap: int):
    """Tile coordinates of predict_with_patches (visualize_vae.py:253-288)."""
    stride = patch_size - overlap
    nh = math.ceil((H - overlap) / stride)
    nw = math.ceil((W - overlap) / stride)
    info = []
    for i in range(nh):
        for j in range(nw):
            sh, sw = i * stride, j * stride
            if i == nh - 1:
                eh = H
                sh = max(0, eh - patch_size)
            else:
                eh = min(sh + patch_size, H)
            if j == nw - 1:
                ew = W
                sw = max(0, ew - patch_size)
            else:
                ew = min(sw + patch_size, W)
            info.append((sh, eh, sw, ew))
    return nh, nw, info


def blend_weight(ph: int, pw: int, i: int, j: int, nh: int, nw: int, overlap: int) -> Tensor:
    """Ramp weights of one tile (visualize_vae.py:361-378): linspace(0,1,overlap) on interior
    edges, only when the tile side exceeds 2*overlap."""
    w = torch.ones(ph, pw)
    ramp = torch.linspace(0, 1, overlap)
    if ph > 2 * overlap:
        if i > 0:
            w[:overlap, :] *= ramp.view(-1, 1)
        if i < nh - 1:
            w[-overlap:, :] *= (1 - ramp).view(-1, 1)
    if pw > 2 * overlap:
        if j > 0:
            w[:, :overlap] *= ramp.view(1, -1)
        if j < nw - 1:
            w[:, -overlap:] *= (1 - ramp).view(1, -1)
    return w


def predict_with_patches(sd: SD, cfg: ResNetCfg, img: Tensor, z: Tensor, patch_size: int = 512,
                         overlap: Optional[int] = None, batch_size: int = 4) -> Tensor:
    """visualize_vae.py:243-415. Per tile: encoder + decode(z) -> sigmoid at tile/2 ->
    bilinear up to the tile -> ramp-weighted accumulate -> divide by (weight + 1e-8).
    (Batching in the reference pads ragged tiles to a common size, :305-318; tiles of one
    image are all equal unless the image is smaller than a patch, so tiles are run one at a
    time here: BN is in eval mode and every op is per-sample, the result is identical.)"""
    B, C, H, W = img.shape
    if overlap is None:
        overlap = max(min(int(patch_size * 0.2), 128), 32)
    nh, nw, info = patch_grid(H, W, patch_size, overlap)
    out = torch.zeros(B, 1, H, W)
    wsum = torch.zeros(B, 1, H, W)
    if z.dim() == 2:
        z = z.unsqueeze(-1).unsqueeze(-1)
    for idx, (sh, eh, sw, ew) in enumerate(info):
        patch = img[:, :, sh:eh, sw:ew]
        feats = resnet34_features(sd, patch)
        pred = torch.sigmoid(decode_from_features(sd, cfg, feats, z, honour_use_skip=False))
        ph, pw = patch.shape[2:]
        if pred.shape[2:] != (ph, pw):
            pred = _up_ac(pred, (ph, pw))
        i, j = divmod(idx, nw)
        wt = blend_weight(ph, pw, i, j, nh, nw, overlap).view(1, 1, ph, pw)
        out[:, :, sh:eh, sw:ew] += pred * wt
        wsum[:, :, sh:eh, sw:ew] += wt
    return out / (wsum + 1e-8)


def segmentation_distribution(sd: SD, cfg: ResNetCfg, img: Tensor, num_samples: int,
                              temperature: float, eps: Tensor, patch_size: Optional[int] = None,
                              overlap: Optional[int] = None):
    """get_segmentation_distribution_from_image, visualize_vae.py:578-652 (tensor part):
    mu/logvar of the whole image; per sample z = mu + eps*exp(.5 lv)*T, then patches or full.
    eps [N,1,L]. Returns ([N,1,H,W] probabilities, mu, logvar)."""
    mu, logvar = unetresnet_encode(sd, img)
    should_sample = cfg.latent_injection not in ("none",)  # :611
    segs = []
    for n in range(num_samples):
        if should_sample:
            z = mu + eps[n] * (torch.exp(0.5 * logvar) * temperature)  # :620-623
        else:
            z = mu
        if patch_size is not None and patch_size > 0:
            ov = overlap if overlap is not None else max(min(int(patch_size * 0.2), 128), 32)
            seg = predict_with_patches(sd, cfg, img, z, patch_size, ov)
        else:
            seg = predict_full_image(sd, cfg, img, z)
        segs.append(seg[0])
    return torch.stack(segs), mu, logvar


def mc_patch_uncertainty(sd: SD, cfg: ResNetCfg, images: Tensor, temperature: float,
                         num_samples: int, eps: Tensor) -> Dict[str, Tensor]:
    """BASELINE config 2 as one call: for each patch run the encoder once, decode N latent
    samples (z = mu + eps[n]*exp(.5 lv)*T, visualize_vae.py:620-623), upsample each logit map to
    the patch size and apply sigmoid (predict_full_image order, visualize_vae.py:81-83), then
    reduce over N with calculate_uncertainty_metrics (visualize_vae.py:90-117).
    eps [N,B,L]. Returns maps [B,H,W] (+ 'samples' [N,B,H,W], 'mu', 'logvar')."""
    feats = resnet34_features(sd, images)
    mu, logvar = latent_heads(sd, feats[-1])
    segs = []
    for n in range(num_samples):
        if cfg.latent_injection != "none":
            z = mu + eps[n] * (torch.exp(0.5 * logvar) * temperature)
        else:
            z = mu
        out = decode_from_features(sd, cfg, feats, z, honour_use_skip=False)
        segs.append(torch.sigmoid(_up_ac(out, images.shape[2:])))
    segs = torch.stack(segs)  # [N,B,1,H,W]
    res = calculate_uncertainty_metrics(segs)
    res["samples"] = segs.squeeze(2)
    res["mu"], res["logvar"] = mu, logvar
    return res


# --------------------------------------------------------------------------------------
# uncertainty reduction + scalar metrics
# --------------------------------------------------------------------------------------
def calculate_uncertainty_metrics(segmentations: Tensor) -> Dict[str, Tensor]:
    """visualize_vae.py:90-117. segmentations [N,B,C,H,W] probabilities. std is UNBIASED."""
    eps = 1e-7
    mean = segmentations.mean(dim=0)
    std = segmentations.std(dim=0)
    entropy = -(mean * torch.log(mean + eps) + (1 - mean) * torch.log(1 - mean + eps))
    se = -(segmentations * torch.log(segmentations + eps)
           + (1 - segmentations) * torch.log(1 - segmentations + eps))
    mi = entropy - se.mean(dim=0)
    cv = std / (mean + eps)
    return {"mean": mean.squeeze(1), "std": std.squeeze(1), "entropy": entropy.squeeze(1),
            "mutual_info": mi.squeeze(1), "coeff_var": cv.squeeze(1)}


def expected_calibration_error(pred: Tensor, gt: Tensor, num_bins: int = 10):
    """utils/uncertainty_metrics.py:108-135. Bins [lo,hi); p == 1.0 falls in no bin."""
    p = pred.flatten()
    g = gt.flatten()
    bounds = torch.linspace(0, 1, num_bins + 1)
    accs = torch.zeros(num_bins)
    confs = torch.zeros(num_bins)
    counts = torch.zeros(num_bins)
    for b in range(num_bins):
        m = (p >= bounds[b]) & (p < bounds[b + 1])
        counts[b] = m.sum()
        if counts[b] > 0:
            accs[b] = g[m].float().mean()
            confs[b] = p[m].mean()
    ece = (counts * (accs - confs).abs()).sum() / counts.sum()
    return ece.item(), accs.numpy(), confs.numpy(), counts.numpy()


def brier_score(pred: Tensor, gt: Tensor) -> float:
    """utils/uncertainty_metrics.py:137-139."""
    return F.mse_loss(pred, gt.float()).item()


def negative_log_likelihood(pred: Tensor, target: Tensor, epsilon: float = 1e-9) -> float:
    """utils/uncertainty_metrics.py:334-338."""
    p = torch.clamp(pred, epsilon, 1 - epsilon)
    return (-(target * torch.log(p) + (1 - target) * torch.log(1 - p))).mean().item()


def uncertainty_error_dice(unc: Tensor, pred_bin: Tensor, target: Tensor, thr: float = 0.2) -> float:
    """utils/uncertainty_metrics.py:340-353."""
    hu = (unc > thr).float()
    em = (pred_bin != target).float()
    inter = (hu * em).sum()
    den = hu.sum() + em.sum()
    if den == 0:
        return 1.0 if inter == 0 else 0.0
    return ((2.0 * inter) / (den + 1e-8)).item()


def threshold_metrics(pred: Tensor, target: Tensor, epsilon: float = 1e-6) -> Dict[str, float]:
    """utils/metrics.py:98-117 (get_all_metrics): everything on (x > 0.5)-thresholded inputs.
    (The reference calls this on raw logits, evaluate.py:76; the quirk lives in the caller.)"""
    p = (pred > 0.5).float()
    t = (target > 0.5).float()
    inter = (p * t).sum()
    den = p.sum() + t.sum()
    dice = 1.0 if den.item() == 0 else ((2.0 * inter + epsilon) / (den + epsilon)).item()
    iou = ((inter + epsilon) / (den - inter + epsilon)).item()
    fp = p.sum() - inter
    fn = t.sum() - inter
    tn = ((1 - p) * (1 - t)).sum()
    return {"dice": dice, "iou": iou,
            "precision": ((inter + epsilon) / (inter + fp + epsilon)).item(),
            "recall": ((inter + epsilon) / (inter + fn + epsilon)).item(),
            "specificity": ((tn + epsilon) / (tn + fp + epsilon)).item(),
            "accuracy": ((p == t).float().sum() / p.numel()).item()}


def evaluate_metrics(sd: SD, cfg: ResNetCfg, batches: Sequence[Tuple[Tensor, Tensor]], eps: Sequence[Tensor],
                     max_samples: Optional[int] = None):
    """evaluate.evaluate, evaluate.py:19-101 (tensor part): eval-mode forward of every (images, masks) batch (the VAE forward
    samples z, unet_resnet.py:211-213: eps[i] is the draw of batch i), `get_all_metrics` on the RAW logits (:76, the
    threshold-logits-at-0.5 quirk), metrics summed and divided by the number of batches (:97), capped by max_samples
    (:27-28,33-34). Predictions of another size than the masks are resized with align_corners=False (:67-73).
    -> (metrics_avg, [sigmoid(logits) of the batches whose samples the reference keeps for plotting, :79-88])."""
    n = len(batches)
    if max_samples is not None and max_samples < n:
        n = max_samples
    sums: Dict[str, float] = {}
    kept = []
    for i, (images, masks) in enumerate(batches):
        if i >= n:
            break
        logits, _, _ = unetresnet_forward(sd, cfg, images, eps[i])      # normalize() has mean 0 / std 1: identity (:15-16)
        if logits.shape != masks.shape:
            logits = F.interpolate(logits, size=masks.shape[2:], mode="bilinear", align_corners=False)
        m = threshold_metrics(logits, masks)
        if len(kept) < 4:
            kept.append(torch.sigmoid(logits))
        for k, v in m.items():
            sums[k] = sums.get(k, 0) + v
    return {k: v / n for k, v in sums.items()}, kept


def generate_ensemble_prediction(sd: SD, cfg: ResNetCfg, image: Tensor, temps: Sequence[float], samples_per_temp: int,
                                 eps: Tensor, weighted: bool = True, patch_size: int = 512, overlap: int = 100) -> Tensor:
    """visualize_vae.generate_ensemble_prediction, visualize_vae.py:1083-1163: weights 1/(|T-1|+0.5) (or 1), normalised
    (:1098-1109); per temperature the mean over samples_per_temp tile-blended probability maps of
    z = mu + eps*exp(.5 logvar)*T (:1131-1152, mu/logvar of the WHOLE image, :1122-1123), weighted sum over temperatures
    (:1155). eps [len(temps)*samples_per_temp, B, L] in the order the reference draws them (temperature-major)."""
    w = torch.tensor([1.0 / (abs(t - 1.0) + 0.5) if weighted else 1.0 for t in temps])
    w = w / w.sum()
    mu, logvar = unetresnet_encode(sd, image)
    should_sample = cfg.latent_injection != "none"                        # :1112
    out = torch.zeros(1, 1, image.shape[2], image.shape[3])
    k = 0
    for ti, t in enumerate(temps):
        acc = torch.zeros_like(out)
        for _ in range(samples_per_temp):
            z = mu + eps[k] * (torch.exp(0.5 * logvar) * t) if should_sample else mu
            k += 1
            acc += predict_with_patches(sd, cfg, image, z, patch_size, overlap)
        out += acc / samples_per_temp * w[ti]
    return out


# --------------------------------------------------------------------------------------
# losses (utils/loss.py)
# --------------------------------------------------------------------------------------
def dice_loss(inputs: Tensor, targets: Tensor, smooth: float = 1.0) -> Tensor:
    """utils/loss.py:6-28."""
    p = torch.sigmoid(inputs)
    if torch.isnan(p).any():
        p = torch.nan_to_num(p, nan=0.0, posinf=1.0, neginf=0.0)
    p = p.reshape(-1)
    t = targets.reshape(-1)
    inter = (p * t).sum()
    ps = torch.clamp(p.sum(), min=smooth / 2.0)
    ts = torch.clamp(t.sum(), min=smooth / 2.0)
    return 1.0 - (2.0 * inter + smooth) / (ps + ts + smooth)


def combined_loss(inputs: Tensor, targets: Tensor, bce_weight: float = 0.5,
                  dice_weight: float = 0.5) -> Tensor:
    """CombinedLoss, utils/loss.py:44-63."""
    bce = F.binary_cross_entropy_with_logits(inputs, targets)
    return bce_weight * bce + dice_weight * dice_loss(inputs, targets)


def ma_focal_loss(inputs: Tensor, targets: Tensor, alpha: float = 0.8, gamma: float = 2.0) -> Tensor:
    """MAFocalLoss, utils/loss.py:66-92 (eps 1e-6 inside the logs, NaN scrubs)."""
    eps = 1e-6
    p = torch.nan_to_num(torch.sigmoid(inputs), nan=0.0, posinf=1.0, neginf=0.0)
    p_t = targets * p + (1 - targets) * (1 - p)
    fw = (1 - p_t).pow(gamma)
    a_t = targets * alpha + (1 - targets) * (1 - alpha)
    bce = -targets * torch.log(p + eps) - (1 - targets) * torch.log(1 - p + eps)
    loss = torch.nan_to_num(a_t * fw * bce, nan=0.0, posinf=1.0, neginf=0.0)
    return loss.mean()


def ma_segmentation_loss(inputs: Tensor, targets: Tensor, dice_weight: float = 0.5,
                         focal_weight: float = 0.5, focal_gamma: float = 2.0,
                         class_weight: float = 0.9) -> Tensor:
    """MASegmentationLoss, utils/loss.py:95-111."""
    return dice_weight * dice_loss(inputs, targets) + \
        focal_weight * ma_focal_loss(inputs, targets, class_weight, focal_gamma)


def kl_with_free_bits(mu: Tensor, logvar: Tensor, free_bits: float = 1e-4) -> Tensor:
    """utils/loss.py:148-170: per-ELEMENT free bits, sum over dims, mean over batch."""
    mu = torch.nan_to_num(mu, nan=0.0)
    logvar = torch.nan_to_num(logvar, nan=0.0)
    kl = 0.5 * (mu.pow(2) + logvar.exp() - logvar - 1)
    kl = torch.clamp(kl, min=-100.0, max=100.0)
    if free_bits > 0:
        kl = torch.max(kl, torch.tensor(free_bits))
    return torch.nan_to_num(kl.sum(dim=1).mean(), nan=1e-8)


def kl_anneal_weight(epoch, kl_start=0.0, kl_end=1.0, warmup_epochs=10, strategy="linear",
                     batch=None, num_batches=None) -> float:
    """KLAnnealer.get_weight, utils/loss.py:114-145."""
    if strategy == "constant":
        return kl_end
    if batch is not None and num_batches is not None:
        progress = (epoch + batch / num_batches) / warmup_epochs
    else:
        progress = epoch / warmup_epochs
    progress = min(progress, 1.0)
    if strategy == "linear":
        return kl_start + progress * (kl_end - kl_start)
    if strategy == "cyclical":
        return kl_start + (progress % 1.0) * (kl_end - kl_start)
    return kl_end


def latent_stats(mu: Tensor, logvar: Tensor) -> Dict[str, float]:
    """calculate_latent_stats, utils/vae_utils.py:79-103."""
    mean_mu = mu.mean(dim=0)
    mean_var = torch.exp(logvar).mean(dim=0)
    active = ((mean_mu.abs() > 0.1) | (mean_var < 0.9) | (mean_var > 1.1)).sum().item()
    kl = 0.5 * (mean_mu.pow(2) + mean_var - logvar.mean(dim=0) - 1)
    return {"active_dims": active, "total_dims": mu.shape[1],
            "activity_ratio": active / mu.shape[1], "total_kl": kl.sum().item(),
            "mean_mu_abs": mean_mu.abs().mean().item(), "mean_var": mean_var.mean().item()}


# --------------------------------------------------------------------------------------
# training step (train.py:385-411)
# --------------------------------------------------------------------------------------
def training_loss(sd: SD, cfg: ResNetCfg, images: Tensor, masks: Tensor, eps: Tensor,
                  beta: float, free_bits: float = 1e-3, lesion: str = "EX",
                  st: Optional[BNState] = None):
    """seg, mu, logvar = model(images) in TRAIN mode; loss = criterion + beta * KL.
    train.py:387-398; criterion per train.py:312-316."""
    if st is None:
        st = BNState(training=True)
    seg, mu, logvar = unetresnet_forward(sd, cfg, images, eps, st)
    recon = ma_segmentation_loss(seg, masks) if lesion == "MA" else combined_loss(seg, masks)
    kl = kl_with_free_bits(mu, logvar, free_bits)
    return recon + beta * kl, recon, kl, seg, mu, logvar, st


def training_grads(sd: SD, cfg: ResNetCfg, images: Tensor, masks: Tensor, eps: Tensor,
                   beta: float, free_bits: float = 1e-3, lesion: str = "EX"):
    """loss.backward() of train.py:403 w.r.t. every floating-point non-buffer entry."""
    leaves = {}
    for k, v in sd.items():
        if v.is_floating_point() and not (k.endswith("running_mean") or k.endswith("running_var")):
            leaves[k] = v.detach().clone().requires_grad_(True)
        else:
            leaves[k] = v
    loss, recon, kl, seg, mu, logvar, st = training_loss(leaves, cfg, images, masks, eps, beta,
                                                        free_bits, lesion)
    names = [k for k, v in leaves.items() if v.requires_grad]
    grads = torch.autograd.grad(loss, [leaves[k] for k in names], allow_unused=True)
    g = {k: (gi if gi is not None else torch.zeros_like(leaves[k])) for k, gi in zip(names, grads)}
    return loss.detach(), recon.detach(), kl.detach(), g, st.updates


def clip_grad_norm_(grads: Dict[str, Tensor], max_norm: float) -> Tensor:
    """torch.nn.utils.clip_grad_norm_(…, 1.0) as called at train.py:408."""
    total = torch.sqrt(sum((g.float() ** 2).sum() for g in grads.values()))
    coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
    for g in grads.values():
        g.mul_(coef)
    return total


def adamw_step(params: Dict[str, Tensor], grads: Dict[str, Tensor], state: Dict[str, dict],
               lr: float = 1e-4, betas=(0.9, 0.999), eps: float = 1e-8, wd: float = 1e-5):
    """optim.AdamW(lr, weight_decay=1e-5) (train.py:322,333), decoupled weight decay."""
    for k, g in grads.items():
        p = params[k]
        s = state.setdefault(k, {"step": 0, "m": torch.zeros_like(p), "v": torch.zeros_like(p)})
        s["step"] += 1
        p.mul_(1 - lr * wd)
        s["m"].mul_(betas[0]).add_(g, alpha=1 - betas[0])
        s["v"].mul_(betas[1]).addcmul_(g, g, value=1 - betas[1])
        bc1 = 1 - betas[0] ** s["step"]
        bc2 = 1 - betas[1] ** s["step"]
        denom = (s["v"].sqrt() / math.sqrt(bc2)).add_(eps)
        p.addcdiv_(s["m"], denom, value=-lr / bc1)


# --------------------------------------------------------------------------------------
# basic attention U-Net (unet/unet_model.py, unet/unet_parts.py)
# --------------------------------------------------------------------------------------
def double_conv(sd: SD, p: str, x: Tensor, st: Optional[BNState] = None) -> Tensor:
    """DoubleConv, unet/unet_parts.py:32-49 (indices 0,1 / 3,4 of the Sequential)."""
    small = sd[p + "double_conv.0.weight"].shape[1] <= 4      # direct fp32 kernel on the network input
    x = _q(F.relu(_bn(sd, p + "double_conv.1.", _conv(sd, p + "double_conv.0.", x, 1, 1, fp32=small), st)))
    x = _q(F.relu(_bn(sd, p + "double_conv.4.", _conv(sd, p + "double_conv.3.", x, 1, 1), st)))
    return x


def unet_up(sd: SD, p: str, x1: Tensor, x2: Tensor, bilinear: bool,
            st: Optional[BNState] = None) -> Tensor:
    """Up.forward, unet/unet_parts.py:82-95: up, pad to skip, gate the skip, cat([skip, up])."""
    if bilinear:
        x1 = _q(F.interpolate(x1, scale_factor=2, mode="bilinear", align_corners=True))
    else:
        x1 = _q(F.conv_transpose2d(x1, _q(sd[p + "up.weight"]), sd[p + "up.bias"], stride=2))
    dy = x2.size(2) - x1.size(2)
    dx = x2.size(3) - x1.size(3)
    x1 = F.pad(x1, [dx // 2, dx - dx // 2, dy // 2, dy - dy // 2])
    x2 = attention_gate(sd, p + "attention.", x1, x2, st)
    return double_conv(sd, p + "conv.", torch.cat([x2, x1], dim=1), st)


def unet_forward(sd: SD, x: Tensor, bilinear: bool = False, st: Optional[BNState] = None) -> Tensor:
    """UNet.forward, unet/unet_model.py:25-36."""
    x1 = double_conv(sd, "inc.", x, st)
    xs = [x1]
    for k in range(1, 5):
        xs.append(double_conv(sd, f"down{k}.maxpool_conv.1.", F.max_pool2d(xs[-1], 2), st))
    y = xs[4]
    for k in range(1, 5):
        y = unet_up(sd, f"up{k}.", y, xs[4 - k], bilinear, st)
    return _conv(sd, "outc.conv.", y, fp32=True)


# --------------------------------------------------------------------------------------
# sample-sharded moments (multi-GPU merge; CPU restatement of vu_moments_partial / vu_moments_finalize,
# which themselves restate calculate_uncertainty_metrics, visualize_vae.py:90-117, in mergeable form)
# --------------------------------------------------------------------------------------
def _bin_entropy(p: Tensor) -> Tensor:
    eps = 1e-7
    return -(p * torch.log(p + eps) + (1 - p) * torch.log(1 - p + eps))


def moments_partial(probs: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """probs [n, ...] -> (mean, M2 = sum (p - mean)^2, sum of per-sample entropies) of this shard."""
    mean = probs.mean(0)
    return mean, ((probs - mean) ** 2).sum(0), _bin_entropy(probs).sum(0)


def moments_finalize(counts: Sequence[float], mean: Tensor, m2: Tensor, sh: Tensor) -> Dict[str, Tensor]:
    """Chan et al. pairwise merge of shard moments [S, ...] -> the five maps of calculate_uncertainty_metrics."""
    n = 0.0
    m = torch.zeros_like(mean[0]); M2 = torch.zeros_like(mean[0]); H = torch.zeros_like(mean[0])
    for k, nb in enumerate(counts):
        if nb <= 0:
            continue
        nn = n + nb
        d = mean[k] - m
        m = m + d * (nb / nn)
        M2 = M2 + m2[k] + d * d * (n * nb / nn)
        H = H + sh[k]
        n = nn
    std = torch.sqrt((M2 / (n - 1)).clamp_min(0)) if n > 1 else torch.full_like(m, float("nan"))
    ent = _bin_entropy(m)
    return {"mean": m, "std": std, "entropy": ent, "mutual_info": ent - H / n, "coeff_var": std / (m + 1e-7)}


# --------------------------------------------------------------------------------------
# sort-based uncertainty metrics - numpy restatements, TEST INFRASTRUCTURE ONLY
# --------------------------------------------------------------------------------------
def sparsification_metrics(pred_probs: Tensor, uncertainties: Tensor, ground_truth: Tensor, num_points: int = 20):
    """utils/uncertainty_metrics.py:175-240: normalised error of the kept pixels when the most uncertain / random pixels are
    removed. The random baseline uses numpy's global RNG exactly as the reference (seed i + 1000 b, choice w/o replacement)."""
    import numpy as np
    B = pred_probs.shape[0]
    eps = 1e-7
    pe = -(ground_truth * torch.log(pred_probs + eps) + (1 - ground_truth) * torch.log(1 - pred_probs + eps))   # :179-181
    pe = pe.reshape(B, -1).detach().cpu().numpy()
    un = uncertainties.reshape(B, -1).detach().cpu().numpy()
    frac = np.linspace(0, 0.99, num_points)
    er, eu = np.zeros(num_points), np.zeros(num_points)
    for b in range(B):
        e, u = pe[b], un[b]
        if np.all(e == e[0]) or np.isnan(e).any():                                  # :194
            continue
        init = e.mean()
        if init <= 0 or np.isnan(init):
            continue
        n = e.shape[0]
        order = np.argsort(u)[::-1]                                                 # :202
        for i, f in enumerate(frac):
            rem = int(n * f)
            keep = n - rem
            if f >= 1.0 or keep <= 0:
                er[i] += 1.0; eu[i] += 1.0
                continue
            np.random.seed(i + b * 1000)                                            # :216
            ridx = np.random.choice(np.arange(n), keep, replace=False)
            er[i] += e[ridx].mean() / init
            eu[i] += e[order[rem:]].mean() / init
    er /= B; eu /= B                                                                # :226-228 (divides by batch_size)
    if er[0] > 0:
        er = er / er[0]
    if eu[0] > 0:
        eu = eu / eu[0]
    for i in range(1, num_points):                                                  # :235-237 monotone envelope
        if eu[i] > eu[i - 1]:
            eu[i] = eu[i - 1]
    return frac, np.nan_to_num(er, nan=1.0, posinf=1.0, neginf=0.0), np.nan_to_num(eu, nan=1.0, posinf=1.0, neginf=0.0)


def uncertainty_error_auc(mean_pred: Tensor, gt_mask: Tensor, uncertainty_map: Tensor):
    """utils/uncertainty_metrics.py:270-283 with sklearn 1.x's roc_auc_score / average_precision_score restated in numpy
    (sklearn/metrics/_ranking.py: _binary_clf_curve - stable sort by score descending, cumulative TP / FP at the last index of
    every run of equal scores; AUROC = trapezoid over (FPR, TPR) from the origin; AP = sum (R_k - R_{k-1}) P_k)."""
    import numpy as np
    pred_binary = (mean_pred > 0.5).float().view(-1).cpu().numpy()
    y = (pred_binary != gt_mask.view(-1).cpu().numpy()).astype(np.int64)
    s = uncertainty_map.view(-1).cpu().numpy()
    if y.min() == y.max():
        return float("nan"), float("nan")                                            # sklearn raises ValueError; :279-281
    order = np.argsort(s, kind="mergesort")[::-1]
    y, s = y[order], s[order]
    ends = np.r_[np.where(np.diff(s))[0], y.size - 1]
    tps = np.cumsum(y)[ends].astype(np.float64)
    fps = 1.0 + ends - tps
    P, N = tps[-1], fps[-1]
    tpr, fpr = np.r_[0.0, tps / P], np.r_[0.0, fps / N]
    auroc = float(np.sum(np.diff(fpr) * (tpr[1:] + tpr[:-1]) * 0.5))
    prec = tps / (tps + fps)
    rec = tps / P
    ap = float(np.sum(np.diff(np.r_[0.0, rec]) * prec))
    return auroc, ap


# --------------------------------------------------------------------------------------
# patch pipeline - numpy restatements, TEST INFRASTRUCTURE ONLY
# --------------------------------------------------------------------------------------
# The reference resizes with PIL (utils/data_loading.py:587-590; `pillow~=10.3.0`, requirements.txt:8 - a third-party
# dependency that is not vendored). Restated from Pillow's src/libImaging/Resample.c (bicubic_filter, precompute_coeffs,
# normalize_coeffs_8bpc, ImagingResampleHorizontal_8bpc / Vertical_8bpc) and Geometry.c (ImagingScaleAffine + nearest);
# tests/test_oracle_golden.py pins it bit for bit against PIL.Image.resize.
def _pil_bicubic(x: float) -> float:
    a = -0.5
    x = -x if x < 0.0 else x
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def _pil_coeffs(in_size: int, out_size: int):
    import math
    import numpy as np
    bits = 32 - 8 - 2
    scale = in_size / out_size
    fscale = scale if scale > 1.0 else 1.0
    support = 2.0 * fscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds, kk = [], np.zeros((out_size, ksize), dtype=np.int64)
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = max(int(center - support + 0.5), 0)
        xmax = min(int(center + support + 0.5), in_size) - xmin
        w = [_pil_bicubic((x + xmin - center + 0.5) / fscale) for x in range(xmax)]
        ww = 0.0
        for v in w:
            ww += v
        for x, v in enumerate(w):
            v = v / ww if ww != 0.0 else v
            kk[xx, x] = int(-0.5 + v * (1 << bits)) if v < 0 else int(0.5 + v * (1 << bits))
        bounds.append((xmin, xmax))
    return bounds, kk, bits


def pil_resize_u8(arr, new_w: int, new_h: int, is_mask: bool):
    """np.array(Image.fromarray(arr).resize((new_w, new_h), NEAREST if is_mask else BICUBIC)) for a uint8 [H,W(,C)] array."""
    import numpy as np
    a = arr if arr.ndim == 3 else arr[..., None]
    H, W, C = a.shape
    if is_mask:
        def tab(n_in, n_out):
            step, xo, t = n_in / n_out, (n_in / n_out) * 0.5, []
            for _ in range(n_out):
                t.append(int(xo)); xo += step
            return np.array(t)
        out = a[tab(H, new_h)][:, tab(W, new_w)]
    else:
        x = a.astype(np.int64)
        if new_w != W:
            bounds, kk, bits = _pil_coeffs(W, new_w)
            o = np.zeros((H, new_w, C), dtype=np.int64)
            for xx, (f, n) in enumerate(bounds):
                o[:, xx] = np.clip(((1 << (bits - 1)) + (x[:, f:f + n] * kk[xx, :n][None, :, None]).sum(1)) >> bits, 0, 255)
            x = o
        if new_h != H:
            bounds, kk, bits = _pil_coeffs(H, new_h)
            o = np.zeros((new_h, x.shape[1], C), dtype=np.int64)
            for yy, (f, n) in enumerate(bounds):
                o[yy] = np.clip(((1 << (bits - 1)) + (x[f:f + n] * kk[yy, :n][:, None, None]).sum(0)) >> bits, 0, 255)
            x = o
        out = x.astype(np.uint8)
    return out if arr.ndim == 3 else out[..., 0]


def patch_pipeline(img_u8, mask_u8, patch_size: int, scale: float, split: str = "train", skip_border_check: bool = False):
    """preprocess + the patch loop of precompute_all_patches (utils/data_loading.py:580-601, 368-398):
    -> (images [T,3,p,p] fp32, masks [T,1,p,p] fp32, coords, has_lesion list) of the kept patches."""
    import numpy as np
    H, W = img_u8.shape[:2]
    new_w, new_h = int(scale * W), int(scale * H)
    img = pil_resize_u8(img_u8, new_w, new_h, False) if (new_w, new_h) != (W, H) else img_u8
    msk = pil_resize_u8(mask_u8, new_w, new_h, True) if (new_w, new_h) != (W, H) else mask_u8
    it = torch.as_tensor((img.astype(np.float32) / 255.0).transpose(2, 0, 1))          # :599-600
    mt = torch.as_tensor((msk > 0).astype(np.float32)).unsqueeze(0)                      # :595
    _, h, w = it.shape
    imgs, masks, coords, lesion = [], [], [], []
    if h < patch_size or w < patch_size:                                                 # :368
        return None, None, coords, lesion
    stride = patch_size // 2
    thr = 0.5 if split == "test" else 0.1                                                # :293-294
    for y in range(0, h - patch_size + 1, stride):
        for x in range(0, w - patch_size + 1, stride):
            ip = it[:, y:y + patch_size, x:x + patch_size]
            mp = mt[:, y:y + patch_size, x:x + patch_size]
            if not skip_border_check:
                black_ratio = (ip.mean(dim=0) < 0.1).float().mean().item()             # :296-299
                if black_ratio > thr:
                    continue
            imgs.append(ip); masks.append(mp); coords.append((y, x)); lesion.append(bool(torch.any(mp > 0.5)))   # :382
    if not imgs:
        return None, None, coords, lesion
    return torch.stack(imgs), torch.stack(masks), coords, lesion
