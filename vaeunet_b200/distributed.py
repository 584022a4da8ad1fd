"""Multi-GPU partitioning of the MC-inference path: one process per GPU, `torch.distributed`
(NCCL over NVLink on the box, gloo in the CPU tests) for the plumbing.

The path shards three ways and only two of them need a collective:

* patch streams (BASELINE config 2 at 2/4/8 GPUs): patches are independent -> `shard_range`, no collective;
* one patch, MC samples sharded (config 4: 15 samples -> 2,2,2,2,2,2,2,1): every rank runs the encoder
  (72 GF, cheaper than moving five feature maps), decodes its own samples and reduces them to per-pixel
  moments (count, mean, M2, sum of sample entropies); ONE all-gather of 3 maps and a Chan merge
  (`vu_moments_finalize`) gives every rank the exact N-sample statistics (unbiased std, visualize_vae.py:96);
* one tiled full image (config 5: 77 tiles -> 10,10,10,10,10,9,9,9): mu/logvar come from a full-image encoder
  pass on rank 0 (broadcast of 64 floats), tiles are dealt round-robin, each rank accumulates the ramp-weighted
  per-sample numerators of ITS tiles; one reduce-scatter by row strips sums them, each rank normalises and
  reduces over N for its strip (blend first, reduce second: every tile of a sample shares z, so overlaps are
  correlated), one all-gather assembles the maps.

The compute kernels are CUDA-only (no CPU fallback). The partitioning / merge logic below is backend
agnostic so `tests/test_distributed_cpu.py` drives it with world_size 2 over gloo with the oracle as the
compute stand-in.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


# ------------------------------------------------------------------ partitioning (pure host logic)
def world(group=None) -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def shard_counts(n: int, world_size: int) -> List[int]:
    """Balanced split: the first n % world ranks get one extra (15 over 8 -> 2,2,2,2,2,2,2,1)."""
    base, extra = divmod(n, world_size)
    return [base + (1 if r < extra else 0) for r in range(world_size)]


def shard_range(n: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous [start, stop) of `n` items for `rank` (patch streams, MC samples)."""
    counts = shard_counts(n, world_size)
    start = sum(counts[:rank])
    return start, start + counts[rank]


def deal_round_robin(n: int, rank: int, world_size: int) -> List[int]:
    """Items rank, rank + world, ... (tiles: neighbouring tiles land on different GPUs, 77 over 8 ->
    10,10,10,10,10,9,9,9)."""
    return list(range(rank, n, world_size))


def strip_rows(H: int, world_size: int) -> int:
    """Rows per rank of the row-strip decomposition (H padded up to a multiple of world)."""
    return (H + world_size - 1) // world_size


# ------------------------------------------------------------------ collectives (backend agnostic)
def _backend(group=None) -> str:
    return dist.get_backend(group) if dist.is_initialized() else "none"


def all_gather_stack(t: torch.Tensor, group=None) -> torch.Tensor:
    """[...] on every rank -> [world, ...] on every rank."""
    _, ws = world(group)
    if ws == 1:
        return t.unsqueeze(0)
    out = torch.empty((ws,) + tuple(t.shape), dtype=t.dtype, device=t.device)
    if _backend(group) == "nccl":
        dist.all_gather_into_tensor(out, t.contiguous(), group=group)
    else:
        parts = [torch.empty_like(t) for _ in range(ws)]
        dist.all_gather(parts, t.contiguous(), group=group)
        out = torch.stack(parts, 0)
    return out


def reduce_scatter_rows(t: torch.Tensor, group=None) -> torch.Tensor:
    """t [world * rows, ...] (same shape on every rank) -> sum over ranks of this rank's row strip [rows, ...]."""
    rank, ws = world(group)
    if ws == 1:
        return t
    assert t.shape[0] % ws == 0
    rows = t.shape[0] // ws
    if _backend(group) == "nccl":
        out = torch.empty((rows,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.reduce_scatter_tensor(out, t.contiguous(), op=dist.ReduceOp.SUM, group=group)
        return out
    buf = t.clone()                      # gloo has no reduce_scatter_tensor: all-reduce, keep the strip
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return buf[rank * rows:(rank + 1) * rows].contiguous()


def broadcast_(t: torch.Tensor, src: int = 0, group=None) -> torch.Tensor:
    if world(group)[1] > 1:
        dist.broadcast(t, src=src, group=group)
    return t


# ------------------------------------------------------------------ sample-sharded MC (config 4)
def merge_sample_moments(count: int, mean: torch.Tensor, m2: torch.Tensor, sum_ent: torch.Tensor,
                         finalize: Callable, group=None, all_counts: Optional[Sequence[int]] = None) -> Dict[str, torch.Tensor]:
    """All-gather per-rank (mean, M2, sum H) - ONE collective - and Chan-merge them. `finalize(counts, mean[S], m2[S],
    sh[S])` is ops.moments_finalize on the GPU (vu_moments_finalize); the CPU tests pass the oracle's restatement.
    all_counts: the per-rank sample counts when every rank can derive them locally (shard_counts(N, world): the usual
    case) - then nothing is exchanged for them and the merge never synchronises with the host. Without it the counts are
    all-gathered and read back (one host sync)."""
    packed = torch.stack([mean, m2, sum_ent], 0)                      # one collective for the three maps
    allp = all_gather_stack(packed, group)                             # [world, 3, ...]
    if all_counts is not None:
        counts = [float(c) for c in all_counts]
        assert len(counts) == allp.shape[0] and counts[world(group)[0]] == float(count)
    else:
        cnt = all_gather_stack(torch.tensor([float(count)], dtype=torch.float32, device=mean.device), group)
        counts = [float(c) for c in cnt.reshape(-1).tolist()]
    keep = [i for i, c in enumerate(counts) if c > 0]                  # ranks without samples contribute nothing
    sel = allp if len(keep) == len(counts) else allp[keep]
    return finalize([counts[i] for i in keep], sel[:, 0].contiguous(), sel[:, 1].contiguous(), sel[:, 2].contiguous())


def mc_predict_sample_sharded(model, images: torch.Tensor, temperature: float, num_samples: int,
                              eps: Optional[torch.Tensor] = None, group=None) -> Dict[str, torch.Tensor]:
    """BASELINE config 4: N MC samples of the same patches split over the ranks of `group`.
    images [B,3,H,W] and eps [N,B,L] are the SAME on every rank (eps is sliced here); every rank returns the
    full-N mean / std / entropy / mutual_info / coeff_var maps [B,H,W]."""
    from . import ops
    rank, ws = world(group)
    eng = model.engine()
    B, _, H, W = images.shape
    if eps is None:   # identical noise on every rank: draw on rank 0, broadcast
        eps = torch.randn(num_samples, B, eng.latent_dim, device=images.device)
        broadcast_(eps, 0, group)
    lo, hi = shard_range(num_samples, rank, ws)
    n_loc = hi - lo
    with torch.no_grad():
        if n_loc > 0:
            lg, mu, lv = eng.mc_logits(images, temperature, n_loc, eps[lo:hi].contiguous(), sample=True)
            probs = ops.resize_logits(lg.reshape(B * n_loc, lg.shape[2], lg.shape[3]), H, W, ops.SIG_POST)
            probs = probs.reshape(B, n_loc, H, W).permute(1, 0, 2, 3).contiguous()      # [n_loc, B, H, W]
            mean, m2, sh = ops.moments_partial(probs)
        else:
            mean = torch.zeros((B, H, W), dtype=torch.float32, device=images.device)
            m2, sh = torch.zeros_like(mean), torch.zeros_like(mean)
        res = merge_sample_moments(n_loc, mean, m2, sh, ops.moments_finalize, group,
                                   all_counts=shard_counts(num_samples, ws))
    return res


# ------------------------------------------------------------------ tiled full image (config 5)
def strip_reduce(acc: torch.Tensor, wsum: torch.Tensor, reduce_fn: Callable, group=None) -> Dict[str, torch.Tensor]:
    """acc [Ns,H,W] = this rank's ramp-weighted per-sample numerators, wsum [H,W] its weight map (0 where it has
    no tile). Reduce-scatter both by row strips, normalise (visualize_vae.py:409), reduce over the Ns samples of
    the strip with `reduce_fn(probs [1,Ns,rows,W]) -> dict of [1,rows,W]`, all-gather the strips.
    -> dict of [H,W] maps, identical on every rank."""
    _, ws = world(group)
    Ns, H, W = acc.shape
    rows = strip_rows(H, ws)
    Hp = rows * ws
    num = torch.zeros((Hp, Ns, W), dtype=acc.dtype, device=acc.device)
    num[:H] = acc.permute(1, 0, 2)
    den = torch.zeros((Hp, W), dtype=acc.dtype, device=acc.device)
    den[:H] = wsum
    num_s = reduce_scatter_rows(num, group)                            # [rows, Ns, W]
    den_s = reduce_scatter_rows(den, group)                            # [rows, W]
    probs = num_s / (den_s + 1e-8).unsqueeze(1)                        # vu_blend_finalize; padded rows: 0 / eps = 0
    maps = reduce_fn(probs.permute(1, 0, 2).unsqueeze(0).contiguous())  # dict of [1, rows, W]
    keys = sorted(maps)
    packed = torch.stack([maps[k].reshape(rows, W) for k in keys], 0)   # [K, rows, W]
    full = all_gather_stack(packed, group)                             # [world, K, rows, W]
    full = full.permute(1, 0, 2, 3).reshape(len(keys), Hp, W)[:, :H]
    return {k: full[i].contiguous() for i, k in enumerate(keys)}


def tiled_image_mc(model, img: torch.Tensor, num_samples: int, patch_size: int = 512, overlap: Optional[int] = None,
                   temperature: float = 1.0, batch_size: int = 4, eps: Optional[torch.Tensor] = None,
                   group=None) -> Dict[str, torch.Tensor]:
    """BASELINE config 5: one full-scale image [1,3,H,W] tiled into patches, N samples, tiles dealt over the ranks.
    Mirrors get_segmentation_distribution_from_image + calculate_uncertainty_metrics
    (visualize_vae.py:578-652, 90-117). -> dict of [H,W] maps on every rank, plus mu / logvar."""
    from . import ops, predict
    rank, ws = world(group)
    eng = model.engine()
    dev = img.device
    L = eng.latent_dim
    with torch.no_grad():
        stats = torch.zeros((2, 1, L), dtype=torch.float32, device=dev)
        if rank == 0:                                   # full-image encoder pass once (1 784 GF at 4288x2848)
            mu, lv = model.encode(img)
            stats[0], stats[1] = mu, lv
        broadcast_(stats, 0, group)
        mu, lv = stats[0], stats[1]
        if eps is None:
            eps = torch.randn(num_samples, 1, L, device=dev)
            broadcast_(eps, 0, group)
        z_rows = ops.reparameterize(mu.contiguous(), lv.contiguous(), eps.float().contiguous(), temperature)[:, 0].contiguous()
        H, W = img.shape[2:]
        ov = predict.default_overlap(patch_size) if overlap is None else overlap
        nh, nw, info = predict.patch_grid(H, W, patch_size, ov)
        mine = deal_round_robin(len(info), rank, ws)
        acc, wsum = predict.blended_samples(model, img, z_rows, patch_size, ov, batch_size, tile_subset=mine,
                                            finalize=False)
        res = strip_reduce(acc, wsum, lambda p: ops.uncertainty_reduce(p, p.shape[2], p.shape[3], input_is_prob=True),
                           group)
    res["mu"], res["logvar"] = mu, lv
    return res


def gradient_group(max_ctas: int = 2):
    """A process group for the data-parallel gradient all-reduces that NCCL may run on at most `max_ctas` CTAs.

    The training step's convolutions are persistent kernels sized to all 148 SMs; an NCCL kernel that takes 8-16 SMs while a
    bucket is reduced turns their single wave into a wave plus a tail. The gradients need ~12 GB/s (122 MB per ~10 ms step) of
    the 900 GB/s NVLink offers, so two CTAs looked plenty - measured, they are not: the
    reductions then run long enough to be exposed at the end of the backward (2 GPUs: 12.01 ms/step on the default communicator,
    12.48 with 2 CTAs, 15.04 with 1), so nothing uses this by default (bench.py: VU_DP_NCCL_CTAS). Falls back to the default group when this torch / NCCL cannot take a
    per-communicator CTA limit."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return None
    try:
        opts = dist.ProcessGroupNCCL.Options()
        opts.config.max_ctas = int(max_ctas)
        opts.config.min_ctas = 1
        return dist.new_group(backend="nccl", pg_options=opts)
    except Exception:
        return dist.group.WORLD
