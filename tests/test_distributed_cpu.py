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
    res = D.merge_sample_moments(hi - lo, mean, m2, sh, O.moments_finalize)
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
