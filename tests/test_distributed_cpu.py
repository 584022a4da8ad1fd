"""world_size-2 gloo tests (CPU) of the multi-GPU host logic in vaeunet_b200/distributed.py: partitioning,
the all-gather + Chan merge of sample-sharded moments, and the reduce-scatter / all-gather row-strip
reduction of a tiled image. The CUDA kernels cannot run here, so the oracle is the compute stand-in; the
SAME distributed functions run over NCCL on the GPU box (scripts/dist_parity.py)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import vaeunet_oracle as O
from vaeunet_b200 import distributed as D


def test_partitions():
    assert D.shard_counts(15, 8) == [2, 2, 2, 2, 2, 2, 2, 1]          # BASELINE config 4
    assert D.shard_counts(77, 8) == [10, 10, 10, 10, 10, 9, 9, 9]    # config 5 tile balance
    for n, w in [(15, 8), (3, 8), (0, 2), (256, 4), (77, 8)]:
        seen = []
        for r in range(w):
            lo, hi = D.shard_range(n, r, w)
            seen += list(range(lo, hi))
        assert seen == list(range(n))
        dealt = sorted(sum((D.deal_round_robin(n, r, w) for r in range(w)), []))
        assert dealt == list(range(n))
    assert D.strip_rows(2848, 8) == 356 and D.strip_rows(701, 8) == 88


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, fn, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ret[rank] = fn(rank, world)
    finally:
        dist.destroy_process_group()


def _run(fn, world=2):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), fn, ret), nprocs=world, join=True)
    return [ret[r] for r in range(world)]


def _sample_case(rank, world):
    g = torch.Generator().manual_seed(5)
    N, B, H, W = 5, 2, 9, 7                                   # 5 samples over 2 ranks -> 3 + 2
    probs = torch.rand(N, B, H, W, generator=g) * 0.98 + 0.01
    lo, hi = D.shard_range(N, rank, world)
    mean, m2, sh = O.moments_partial(probs[lo:hi])
    res = D.merge_sample_moments(hi - lo, mean, m2, sh, O.moments_finalize)
    ref = O.calculate_uncertainty_metrics(probs.unsqueeze(2))
    # the sync-free form: every rank derives the per-rank counts locally, ONE all-gather, no host read-back
    res2 = D.merge_sample_moments(hi - lo, mean, m2, sh, O.moments_finalize, all_counts=D.shard_counts(N, world))
    assert all(torch.equal(res[k], res2[k]) for k in res)
    return {k: float((res[k] - ref[k]).abs().max()) for k in ref}


def test_sample_sharded_moments_merge_gloo():
    for errs in _run(_sample_case):
        for k, e in errs.items():
            assert e < 2e-6, (k, e)


def _empty_shard_case(rank, world):
    g = torch.Generator().manual_seed(6)
    probs = torch.rand(1, 1, 4, 4, generator=g)               # 1 sample over 2 ranks: rank 1 holds nothing
    lo, hi = D.shard_range(1, rank, world)
    if hi > lo:
        mean, m2, sh = O.moments_partial(probs[lo:hi])
    else:
        mean = torch.zeros(1, 4, 4); m2 = torch.zeros(1, 4, 4); sh = torch.zeros(1, 4, 4)
    res = D.merge_sample_moments(hi - lo, mean, m2, sh, O.moments_finalize, all_counts=D.shard_counts(1, world))
    return float((res["mean"] - probs[0]).abs().max()), bool(torch.isnan(res["std"]).all())


def test_empty_shard_gloo():
    for err, std_nan in _run(_empty_shard_case):
        assert err < 1e-7 and std_nan                         # torch.std of one sample is NaN (visualize_vae.py:96)


def _tile_case(rank, world):
    g = torch.Generator().manual_seed(7)
    Ns, H, W, P, ov = 3, 37, 53, 24, 6                        # ragged grid, H not a multiple of world
    nh, nw, info = O.patch_grid(H, W, P, ov)
    tiles = [torch.rand(Ns, eh - sh, ew - sw, generator=g) for (sh, eh, sw, ew) in info]

    def accumulate(idxs):
        acc = torch.zeros(Ns, H, W); wsum = torch.zeros(H, W)
        for k in idxs:
            sh, eh, sw, ew = info[k]
            w = O.blend_weight(eh - sh, ew - sw, k // nw, k % nw, nh, nw, ov)
            acc[:, sh:eh, sw:ew] += tiles[k] * w
            wsum[sh:eh, sw:ew] += w
        return acc, wsum

    def reduce_fn(p):                                          # p [1, Ns, rows, W] -> dict of [1, rows, W]
        return O.calculate_uncertainty_metrics(p.permute(1, 0, 2, 3).unsqueeze(2))

    acc, wsum = accumulate(D.deal_round_robin(len(info), rank, world))
    res = D.strip_reduce(acc, wsum, reduce_fn)
    acc_all, wsum_all = accumulate(range(len(info)))
    ref = O.calculate_uncertainty_metrics((acc_all / (wsum_all + 1e-8)).unsqueeze(1).unsqueeze(1))
    return {k: float((res[k] - ref[k][0]).abs().max()) for k in ref}, tuple(res["mean"].shape)


def test_tiled_strip_reduce_gloo():
    for errs, shape in _run(_tile_case):
        assert shape == (37, 53)
        for k, e in errs.items():
            assert e < 5e-6, (k, e)


# ---------------------------------------------------------------------------------------------------------
# data-parallel training: flat-buffer layout, bucketed all-reduce, averaged AdamW step
# ---------------------------------------------------------------------------------------------------------
def _layout():
    from vaeunet_b200 import testing
    from vaeunet_b200.training import BUCKET_ORDER, param_layout
    sd, _ = testing.resnet_fixture()
    named = [(k, tuple(v.shape)) for k, v in sd.items()
             if v.is_floating_point() and not k.endswith(("running_mean", "running_var"))]
    return sd, named, param_layout(named), BUCKET_ORDER


def test_param_layout_buckets():
    sd, named, (offs, total, buckets), order = _layout()
    assert set(buckets) == set(order)
    spans = sorted(buckets.values())
    assert spans[0][0] == 0 and spans[-1][1] == total
    for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
        assert a1 == b0, "buckets must tile the flat buffer without gaps or overlap"
    for n, shp in named:
        k = 1
        for d in shp:
            k *= d
        assert offs[n] % 64 == 0                                          # 256-byte aligned
        inside = [b for b, (lo, hi) in buckets.items() if lo <= offs[n] and offs[n] + k <= hi]
        assert len(inside) == 1, n
    assert len(named) == 206 and total >= 30_455_789                    # 30.46 M parameters in 206 tensors


def _dp_case(rank, world):
    """Each rank: oracle gradients of its own batch -> flat buffer -> bucketed async all-reduce in backward order ->
    AdamW on grad / world. Returns the updated flat parameters and the reference computed from gathered gradients."""
    from vaeunet_b200.synthetic import synthetic_fundus
    sd, named, (offs, total, buckets), order = _layout()
    torch.set_num_threads(2)
    x, m = synthetic_fundus(2, 64, 64, seed=50 + rank)
    eps = torch.randn(2, 32, generator=torch.Generator().manual_seed(60 + rank))
    _, _, _, grads, _ = O.training_grads(sd, O.ResNetCfg(), x, m, eps, 0.001, 1e-3, "EX")
    flat = torch.zeros(total)
    for n, shp in named:
        flat[offs[n]:offs[n] + grads[n].numel()] = grads[n].flatten()
    local = flat.clone()
    pending = [dist.all_reduce(flat[buckets[b][0]:buckets[b][1]], async_op=True) for b in order]
    for w in pending:
        w.wait()
    gathered = [torch.zeros(total) for _ in range(world)]
    dist.all_gather(gathered, local)
    want = sum(gathered)
    err_sum = float((flat - want).abs().max())

    def step(gflat):
        g = {n: (gflat[offs[n]:offs[n] + sd[n].numel()] / world).view(sd[n].shape).clone() for n, _ in named}
        p = {n: sd[n].clone() for n, _ in named}
        O.clip_grad_norm_(g, 1.0)
        O.adamw_step(p, g, {}, lr=1e-4, wd=1e-5)
        return torch.cat([p[n].flatten() for n, _ in named])
    mine = step(flat)
    p0 = mine.clone()
    dist.broadcast(p0, 0)
    return {"err_sum": err_sum, "same_as_rank0": float((mine - p0).abs().max()),
            "moved": float((mine - torch.cat([sd[n].flatten() for n, _ in named])).abs().max())}


def test_dp_training_step_gloo():
    res = _run(_dp_case, 2)
    for r in res:
        assert r["err_sum"] < 1e-5          # bucketed all-reduce == sum of the ranks' local gradients
        assert r["same_as_rank0"] == 0.0    # identical parameters on every rank after the step
        assert 0 < r["moved"] <= 1.1e-4     # first AdamW step moves each weight by at most lr (+ decay)
