"""Drop-ins for the hot-path functions of the reference's visualize_vae.py:
predict_full_image (:61-87), calculate_uncertainty_metrics (:90-117), predict_with_patches
(:243-415) and the N-sample loop of get_segmentation_distribution_from_image (:578-652).

B200 re-design: tiles are batched through the engine; blending, normalisation and the
N-sample reduction stay on the device (the reference round-trips every tile through the CPU,
:380-384, and re-runs the encoder for every sample, :324 — `segmentation_distribution` below runs
the encoder ONCE per tile batch and decodes all N samples as one batched pass)."""
from __future__ import annotations

import math
from typing import Sequence, Dict, Optional

import torch

from . import ops


def _z_vec(z: torch.Tensor) -> torch.Tensor:
    return z.float().reshape(z.shape[0], z.shape[1]).contiguous()


def predict_full_image(model, img, z):
    """visualize_vae.py:61-87: encoder + decode one z -> upsample -> sigmoid. z [1,L,1,1] or [1,L]."""
    model.eval()
    eng = model.engine()
    with torch.no_grad():
        feats = eng.encode(img.float().contiguous())
        lg = eng.decode(feats, _z_vec(z).expand(img.shape[0], -1).contiguous(), honour_use_skip=False)
        B, C, h, w = lg.shape
        out = ops.resize_logits(lg.reshape(B * C, h, w), img.shape[2], img.shape[3], ops.SIG_POST)
        return out.reshape(B, C, img.shape[2], img.shape[3])


def calculate_uncertainty_metrics(segmentations: torch.Tensor) -> Dict[str, torch.Tensor]:
    """visualize_vae.py:90-117. segmentations [N,B,1,H,W] (or [N,1,H,W]) probabilities on the device."""
    s = segmentations.float()
    if s.dim() == 4:
        s = s.unsqueeze(1)
    N, B, C, H, W = s.shape
    assert C == 1
    x = s[:, :, 0].permute(1, 0, 2, 3).contiguous()   # [B,N,H,W]
    return ops.uncertainty_reduce(x, H, W, input_is_prob=True)


def patch_grid(H: int, W: int, patch_size: int, overlap: int):
    """Tile coordinates, visualize_vae.py:253-288 (last row/col snapped to the image edge)."""
    stride = patch_size - overlap
    nh = math.ceil((H - overlap) / stride)
    nw = math.ceil((W - overlap) / stride)
    info = []
    for i in range(nh):
        for j in range(nw):
            sh, sw = i * stride, j * stride
            if i == nh - 1:
                eh = H; sh = max(0, eh - patch_size)
            else:
                eh = min(sh + patch_size, H)
            if j == nw - 1:
                ew = W; sw = max(0, ew - patch_size)
            else:
                ew = min(sw + patch_size, W)
            info.append((sh, eh, sw, ew))
    return nh, nw, info


def default_overlap(patch_size: int) -> int:
    return max(min(int(patch_size * 0.2), 128), 32)   # visualize_vae.py:251


def _tiles_probs(eng, img, z_dec_rows: torch.Tensor, info, idxs, Ns: int):
    """Run a batch of equal-sized tiles: ONE gather kernel cuts them out of the device-resident image (no torch.cat of
    slices), encoder once, Ns decodes each (sample-minor).
    -> probabilities [T, Ns, ph, pw] (sigmoid at tile/2, THEN bilinear up: visualize_vae.py:345-358)."""
    ph, pw = info[idxs[0]][1] - info[idxs[0]][0], info[idxs[0]][3] - info[idxs[0]][2]
    coords = torch.tensor([[info[k][0], info[k][2]] for k in idxs], dtype=torch.int32).to(img.device, non_blocking=True)
    tiles = ops.tile_gather(img, coords, ph, pw)
    T = tiles.shape[0]
    feats = eng.encode(tiles)
    z_dec = z_dec_rows.unsqueeze(0).expand(T, -1, -1).reshape(T * Ns, -1).contiguous()
    lg = eng.decode(feats, z_dec, honour_use_skip=False)          # [T*Ns,1,h,w]
    h, w = lg.shape[2:]
    return ops.resize_logits(lg.reshape(T * Ns, h, w), ph, pw, ops.SIG_PRE).reshape(T, Ns, ph, pw)


def _blend_weight_map(ph: int, pw: int, i: int, j: int, nh: int, nw: int, overlap: int, device) -> torch.Tensor:
    """Ramp weights of tile (i, j) in an nh x nw grid (visualize_vae.py:361-378 == :451-470): linspace(0, 1, overlap) on
    every edge that has a neighbour, only when the tile is longer than 2 * overlap along that axis. Built by running the
    blend kernel on a tile of ones into an empty canvas of the tile's own size."""
    ones = torch.ones((ph, pw), dtype=torch.float32, device=device)
    acc = torch.zeros((ph, pw), dtype=torch.float32, device=device)
    ws = torch.zeros((ph, pw), dtype=torch.float32, device=device)
    ops.blend_accumulate(ones, 0, 0, i, j, nh, nw, overlap, acc, ws)
    return ws


def predict_single_patch(model, patch, z, device=None, overlap=100, patch_grid=(1, 1), patch_idx=0):
    """visualize_vae.py:418-476 (the memory-fallback helper of predict_with_patches): one tile through encoder + decode(z)
    -> sigmoid at tile/2 -> bilinear up to the tile, plus the tile's ramp blend weight. -> (patch_pred [B,1,h,w], weight
    [B,1,h,w]). (The reference's version always starts from z_initial whatever `use_bottleneck` says; models
    without a bottleneck latent go through predict_with_patches' own path.)"""
    model.eval()
    eng = model.engine()
    nh, nw = patch_grid
    i, j = divmod(int(patch_idx), nw)
    with torch.no_grad():
        patch = patch.float().contiguous()
        B, _, ph, pw = patch.shape
        feats = eng.encode(patch)
        lg = eng.decode(feats, _z_vec(z).expand(B, -1).contiguous(), honour_use_skip=False)
        h, w = lg.shape[2:]
        pred = ops.resize_logits(lg.reshape(B, h, w), ph, pw, ops.SIG_PRE).reshape(B, 1, ph, pw)
        wt = _blend_weight_map(ph, pw, i, j, nh, nw, int(overlap), patch.device)
    return pred, wt.view(1, 1, ph, pw).expand(B, 1, ph, pw).contiguous()


def predict_with_patches(model, img, z, patch_size=512, overlap=None, batch_size=4):
    """visualize_vae.py:243-415 for one z ([1,L,1,1]). -> [1,1,H,W] blended probabilities."""
    out = blended_samples(model, img, _z_vec(z), patch_size, overlap, batch_size)
    return out[:, None]


def blended_samples(model, img, z_rows: torch.Tensor, patch_size=512, overlap=None, batch_size=4,
                    tile_subset=None, finalize=True):
    """All Ns samples of a tiled image in one sweep: z_rows [Ns, L] (every tile of a sample shares
    its z, visualize_vae.py:328-334). -> [Ns,H,W] blended probabilities (or the raw accumulators
    (acc [Ns,H,W], wsum [H,W]) when finalize=False, for multi-GPU tile sharding).
    Per tile batch: one gather launch, the network, ONE blend launch for all tiles x samples; the ramp weights do not depend
    on the sample, so there is a single weight map."""
    model.eval()
    eng = model.engine()
    B, C, H, W = img.shape
    assert B == 1, "the reference tiles one image at a time"
    if overlap is None:
        overlap = default_overlap(patch_size)
    nh, nw, info = patch_grid(H, W, patch_size, overlap)
    Ns = z_rows.shape[0]
    img = img.float().contiguous()
    acc = torch.zeros((Ns, H, W), dtype=torch.float32, device=img.device)
    wsum = torch.zeros((H, W), dtype=torch.float32, device=img.device)
    todo = list(range(len(info))) if tile_subset is None else list(tile_subset)
    with torch.no_grad():
        # group equal-sized tiles (ragged only when the image is smaller than a patch)
        pos = 0
        while pos < len(todo):
            grp = [todo[pos]]
            shp = (info[todo[pos]][1] - info[todo[pos]][0], info[todo[pos]][3] - info[todo[pos]][2])
            pos += 1
            while pos < len(todo) and len(grp) < batch_size and \
                    (info[todo[pos]][1] - info[todo[pos]][0], info[todo[pos]][3] - info[todo[pos]][2]) == shp:
                grp.append(todo[pos]); pos += 1
            probs = _tiles_probs(eng, img, z_rows, info, grp, Ns)
            rows = [[info[k][0], info[k][2], k // nw, k % nw] for k in grp]
            tinfo = torch.tensor(rows, dtype=torch.int32).to(img.device, non_blocking=True)
            bbox = (min(info[k][0] for k in grp), max(info[k][1] for k in grp),
                    min(info[k][2] for k in grp), max(info[k][3] for k in grp))
            ops.blend_accumulate_batch(probs, tinfo, bbox, nh, nw, overlap, acc, wsum)
    if not finalize:
        return acc, wsum
    return ops.blend_finalize_shared(acc, wsum)


def segmentation_distribution(model, img, num_samples=32, patch_size=None, overlap=None, temperature=1.0,
                              batch_size=4, eps: Optional[torch.Tensor] = None):
    """Tensor part of get_segmentation_distribution_from_image (visualize_vae.py:578-652):
    mu/logvar from the WHOLE image (:603), N samples z = mu + eps*exp(.5 logvar)*T (:620-623), each
    decoded by patches or full image. -> (segmentations [N,1,H,W] on the device, mu, logvar)."""
    model.eval()
    with torch.no_grad():
        mu, logvar = model.encode(img)
        sample = getattr(model, "latent_injection", "all") not in ("none",)
        if sample:
            if eps is None:
                eps = torch.randn(num_samples, *mu.shape, device=mu.device)
            z = ops.reparameterize(mu, logvar, eps.float().contiguous(), temperature)   # [N,1,L]
        else:
            z = mu.unsqueeze(0).expand(num_samples, -1, -1).contiguous()
        z_rows = z[:, 0].contiguous()
        if patch_size is not None and patch_size > 0:
            segs = blended_samples(model, img, z_rows, patch_size, overlap, batch_size)
        else:
            eng = model.engine()
            feats = eng.encode(img.float().contiguous())
            lg = eng.decode(feats, z_rows, honour_use_skip=False)
            segs = ops.resize_logits(lg.reshape(num_samples, lg.shape[2], lg.shape[3]), img.shape[2], img.shape[3],
                                     ops.SIG_POST)
    return segs.unsqueeze(1), mu, logvar


def ensemble_weights(temps: Sequence[float], weighted: bool = True) -> torch.Tensor:
    """visualize_vae.py:1098-1109: 1 / (|T - 1| + 0.5) per temperature (or 1), normalised to sum 1."""
    w = torch.tensor([1.0 / (abs(float(t) - 1.0) + 0.5) if weighted else 1.0 for t in temps], dtype=torch.float32)
    return w / w.sum()


def generate_ensemble_prediction(model, image, temps=(0.5, 1.0, 2.0, 3.0), samples_per_temp=5, weighted=True, patch_size=512,
                                 overlap=100, batch_size=4, eps: Optional[torch.Tensor] = None) -> torch.Tensor:
    """visualize_vae.generate_ensemble_prediction (visualize_vae.py:1083-1163): temperature-weighted mean of tile-blended
    probability maps, samples_per_temp draws z = mu + eps*exp(.5 logvar)*T per temperature. -> [1,1,H,W].

    The reference re-encodes the image per temperature and runs len(temps)*samples_per_temp sequential tiled predictions, each
    re-running the encoder on every tile. Here mu/logvar are computed once and ALL draws go through one tiled sweep
    (`blended_samples`): the encoder runs once per tile and the decodes of every (T, sample) pair share its features.
    eps (additive): [len(temps)*samples_per_temp, 1, L], temperature-major (the reference's draw order)."""
    model.eval()
    T, S = len(temps), int(samples_per_temp)
    with torch.no_grad():
        mu, logvar = model.encode(image)
        if getattr(model, "latent_injection", "all") != "none":                       # :1112
            if eps is None:
                eps = torch.randn(T * S, *mu.shape, device=mu.device)
            eps = eps.float().contiguous().to(mu.device)
            z = torch.cat([ops.reparameterize(mu, logvar, eps[t * S:(t + 1) * S].contiguous(), float(temps[t]))
                           for t in range(T)])                                        # [T*S, 1, L]
        else:
            z = mu.unsqueeze(0).expand(T * S, -1, -1).contiguous()
        segs = blended_samples(model, image, z[:, 0].contiguous(), patch_size, overlap, batch_size)     # [T*S, H, W]
        w = (ensemble_weights(temps, weighted) / S).to(segs.device).repeat_interleave(S)                # weight of every draw
        return (segs * w.view(-1, 1, 1)).sum(0)[None, None]
