"""GPU parity of the memory-bound kernels (through the C ABI) against torch CPU fp32 primitives,
i.e. the same calls the oracle makes."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _q(x):  # bf16 round trip
    return x.bfloat16().float()


def _to_act(ops, x_nchw, dev):
    return ops.Act(x_nchw.permute(0, 2, 3, 1).contiguous().bfloat16().to(dev))


def _from_act(a):
    return a.t[..., a.coff:a.coff + a.C].float().cpu().permute(0, 3, 1, 2)


def test_layout_roundtrip(cuda_dev):
    from vaeunet_b200 import ops
    x = torch.randn(2, 37, 19, 23)
    a = ops.nchw_to_nhwc(x.to(cuda_dev))
    assert torch.equal(_from_act(a), _q(x))
    assert torch.equal(ops.nhwc_to_nchw(a).cpu(), _q(x))


@pytest.mark.parametrize("H,W", [(64, 64), (50, 70)])
def test_stem_conv(cuda_dev, H, W):
    from vaeunet_b200 import ops
    g = torch.Generator().manual_seed(1)
    x = torch.rand(2, 3, H, W, generator=g)
    w = torch.randn(64, 3, 7, 7, generator=g) * 0.1
    s = torch.rand(64, generator=g) + 0.5
    b = torch.randn(64, generator=g) * 0.1
    ref = F.relu(F.conv2d(x, w, None, 2, 3) * s.view(1, -1, 1, 1) + b.view(1, -1, 1, 1))
    got = _from_act(ops.stem_conv7x7s2(x.to(cuda_dev), w.to(cuda_dev), s.to(cuda_dev), b.to(cuda_dev)))
    assert (got - ref).abs().max() <= 1e-2 * ref.abs().max()


def test_conv3x3_small_cin(cuda_dev):
    from vaeunet_b200 import ops
    g = torch.Generator().manual_seed(2)
    x = torch.rand(2, 3, 40, 33, generator=g)
    w = torch.randn(64, 3, 3, 3, generator=g) * 0.2
    s = torch.rand(64, generator=g) + 0.5
    b = torch.randn(64, generator=g) * 0.1
    ref = F.relu(F.conv2d(x, w, None, 1, 1) * s.view(1, -1, 1, 1) + b.view(1, -1, 1, 1))
    got = _from_act(ops.conv3x3_small_cin(x.to(cuda_dev), w.to(cuda_dev), s.to(cuda_dev), b.to(cuda_dev), True))
    assert (got - ref).abs().max() <= 1e-2 * ref.abs().max()


@pytest.mark.parametrize("k,s,p,H,W", [(3, 2, 1, 32, 32), (3, 2, 1, 35, 29), (2, 2, 0, 32, 30)])
def test_maxpool(cuda_dev, k, s, p, H, W):
    from vaeunet_b200 import ops
    x = _q(torch.randn(2, 64, H, W))
    ref = F.max_pool2d(x, k, s, p)
    got = _from_act(ops.maxpool(_to_act(ops, x, cuda_dev), k, s, p))
    assert torch.equal(got, ref)


@pytest.mark.parametrize("C", [32, 24, 128, 512])       # 24: generic kernel; powers of two: the pixel-run kernel
@pytest.mark.parametrize("Hs,Ws,Ho,Wo", [(16, 16, 32, 32), (22, 22, 44, 44), (88, 88, 175, 175), (1, 1, 8, 8), (5, 7, 5, 7)])
def test_bilinear(cuda_dev, Hs, Ws, Ho, Wo, C):
    from vaeunet_b200 import ops
    x = _q(torch.randn(2, C, Hs, Ws))
    ref = F.interpolate(x, size=(Ho, Wo), mode="bilinear", align_corners=True)
    got = _from_act(ops.bilinear_ac(_to_act(ops, x, cuda_dev), Ho, Wo))
    assert (got - ref).abs().max() <= 8e-3 * max(ref.abs().max().item(), 1.0)


def test_bilinear_padded_slice(cuda_dev):
    from vaeunet_b200 import ops
    x = _q(torch.randn(1, 16, 8, 8))
    out = ops.Act(torch.zeros(1, 19, 18, 48, dtype=torch.bfloat16, device=cuda_dev), 16, 16)
    ops.bilinear_ac(_to_act(ops, x, cuda_dev), 16, 16, out=out, pad_t=1, pad_l=1)
    ref = F.pad(F.interpolate(x, size=(16, 16), mode="bilinear", align_corners=True), [1, 1, 1, 2])
    assert (_from_act(out) - ref).abs().max() <= 8e-3 * ref.abs().max()
    assert (out.t[..., :16] == 0).all() and (out.t[..., 32:] == 0).all()


def test_broadcast_vec(cuda_dev):
    from vaeunet_b200 import ops
    v = torch.randn(3, 32)
    out = ops.new_act(3, 5, 7, 32)
    ops.broadcast_vec(v.to(cuda_dev), out)
    assert torch.equal(_from_act(out), _q(v).view(3, 32, 1, 1).expand(3, 32, 5, 7))


@pytest.mark.parametrize("F_,Cl,rep", [(128, 256, 3), (32, 64, 2), (256, 512, 1), (64, 64, 1)])
def test_attention_gate(cuda_dev, F_, Cl, rep):
    from vaeunet_b200 import ops
    g = torch.Generator().manual_seed(3)
    Ns, H, W = 2, 9, 11
    Ng = Ns * rep
    g1 = _q(torch.randn(Ng, F_, H, W, generator=g))
    x1 = _q(torch.randn(Ns, F_, H, W, generator=g))
    skip = _q(torch.randn(Ns, Cl, H, W, generator=g))
    pw = torch.randn(F_, generator=g) * 0.2
    ps, pb = 0.7, -0.1
    x1r = x1.repeat_interleave(rep, 0)
    skr = skip.repeat_interleave(rep, 0)
    psi = torch.sigmoid(ps * (F.relu(g1 + x1r) * pw.view(1, -1, 1, 1)).sum(1, keepdim=True) + pb)
    ref = skr * psi
    out = ops.new_act(Ng, H, W, Cl)
    psi_out = torch.empty(Ng, H, W, device=cuda_dev)
    ops.attention_gate(_to_act(ops, g1, cuda_dev), _to_act(ops, x1, cuda_dev), _to_act(ops, skip, cuda_dev),
                       pw.to(cuda_dev), ps, pb, out, psi_out)
    assert (psi_out.cpu() - psi[:, 0]).abs().max() < 1e-4
    assert (_from_act(out) - ref).abs().max() <= 8e-3 * ref.abs().max()


@pytest.mark.parametrize("hs,ws,H,W,Cx,Cs,F_,Cz,rep", [(8, 8, 16, 16, 128, 64, 32, 32, 2), (11, 9, 22, 17, 64, 128, 64, 0, 1),
                                                        (1, 1, 6, 5, 256, 64, 128, 32, 3), (5, 7, 10, 14, 24, 8, 8, 8, 1),
                                                        # the fp16 fast path (Cz = 0, Cx == Cs in {64,128,256}): ragged rows, several runs per row
                                                        (8, 8, 16, 16, 64, 64, 32, 0, 3), (11, 21, 22, 41, 128, 128, 64, 0, 2),
                                                        (4, 4, 8, 8, 256, 256, 128, 0, 2), (16, 20, 32, 40, 64, 64, 32, 0, 2),
                                                        (1, 1, 4, 36, 128, 128, 64, 0, 2),
                                                        # the decoder blocks' own slice widths (unet_resnet.py:228-231)
                                                        (8, 8, 16, 16, 128, 64, 32, 0, 3), (9, 11, 18, 21, 128, 64, 32, 0, 2),
                                                        (5, 6, 10, 12, 256, 64, 64, 0, 2), (3, 3, 6, 7, 512, 128, 128, 0, 2),
                                                        (2, 3, 4, 6, 512, 256, 128, 0, 2)])
@pytest.mark.parametrize("dt", ["bf16", "f16"])
def test_decoder_concat(cuda_dev, hs, ws, H, W, Cx, Cs, F_, Cz, rep, dt):
    """Fused concat producer vs. the reference order of operations (unet/unet_resnet.py:79-98): upsample x to the
    skip size, g1 = affine(W_g(x_up)) at OUTPUT resolution, gate, concat [x_up | skip*psi | z]."""
    from vaeunet_b200 import ops
    g = torch.Generator().manual_seed(31)
    tdt = torch.bfloat16 if dt == "bf16" else torch.float16
    odt = ops.BF16 if dt == "bf16" else ops.F16
    _q = lambda t: t.to(tdt).float()                                                   # noqa: E731
    _to_act = lambda o, t, dv: o.Act(t.permute(0, 2, 3, 1).contiguous().to(tdt).to(dv))  # noqa: E731
    Dp = 2
    D = Dp * rep
    x = _q(torch.randn(D, Cx, hs, ws, generator=g))
    wg = torch.randn(F_, Cx, 1, 1, generator=g) * (1.0 / Cx ** 0.5)
    bg = torch.randn(F_, generator=g) * 0.1
    g1_low = _q(F.conv2d(x, wg, bg))                      # what the tensor-core conv stores at low resolution
    x1 = _q(torch.randn(Dp, F_, H, W, generator=g))
    skip = _q(torch.randn(Dp, Cs, H, W, generator=g))
    pw = torch.randn(F_, generator=g) * 0.2
    ps, pb = 0.8, 0.05
    zv = torch.randn(D, Cz, generator=g) if Cz else None
    x_up = F.interpolate(x, size=(H, W), mode="bilinear", align_corners=True)
    g1_ref = F.conv2d(x_up, wg, bg)                       # reference order: conv after the upsample
    psi = torch.sigmoid(ps * (F.relu(g1_ref + x1.repeat_interleave(rep, 0)) * pw.view(1, -1, 1, 1)).sum(1, keepdim=True) + pb)
    parts = [x_up, skip.repeat_interleave(rep, 0) * psi]
    if Cz:
        parts.append(zv.view(D, Cz, 1, 1).expand(D, Cz, H, W))
    ref = torch.cat(parts, 1)
    d = cuda_dev
    out = ops.new_act(D, H, W, Cx + Cs + Cz, odt)
    psi_out = torch.empty(D, H, W, device=d)
    ops.decoder_concat(_to_act(ops, x, d), _to_act(ops, g1_low, d), _to_act(ops, x1, d), _to_act(ops, skip, d),
                       pw.to(d), ps, pb, zv.to(d) if Cz else None, out, psi_out)
    assert (psi_out.cpu() - psi[:, 0]).abs().max() < 5e-3          # g1 stored in bf16 at low resolution
    got = _from_act(out)
    assert (got[:, :Cx] - x_up).abs().max() <= 8e-3 * x_up.abs().max()
    assert (got - ref).abs().max() <= 1.2e-2 * ref.abs().max()


def test_latent_head_and_project(cuda_dev):
    from vaeunet_b200 import ops
    g = torch.Generator().manual_seed(4)
    B, Cc, h, w, L, Ns = 3, 512, 7, 5, 32, 4
    f4 = _q(torch.randn(B, Cc, h, w, generator=g))
    wm = torch.randn(L, Cc, generator=g) * 0.05
    bm = torch.randn(L, generator=g) * 0.1
    wl = torch.randn(L, Cc, generator=g) * 0.05
    bl = torch.randn(L, generator=g) * 0.1
    eps = torch.randn(Ns, B, L, generator=g)
    T = 1.7
    mu_r = F.adaptive_avg_pool2d(F.conv2d(f4, wm.view(L, Cc, 1, 1), bm), 1).flatten(1)
    lv_r = F.adaptive_avg_pool2d(F.conv2d(f4, wl.view(L, Cc, 1, 1), bl), 1).flatten(1)
    z_r = mu_r + eps * (torch.exp(0.5 * lv_r) * T)
    d = cuda_dev
    mu, lv, z = ops.latent_head(_to_act(ops, f4, d), wm.to(d), bm.to(d), wl.to(d), bl.to(d), eps.to(d), T, Ns)
    assert (mu.cpu() - mu_r).abs().max() < 1e-4 and (lv.cpu() - lv_r).abs().max() < 1e-4
    assert (z.cpu() - z_r).abs().max() < 2e-4
    _, _, z0 = ops.latent_head(_to_act(ops, f4, d), wm.to(d), bm.to(d), wl.to(d), bl.to(d), None, T, 1)
    assert (z0.cpu()[0] - mu_r).abs().max() < 1e-4
    wz = torch.randn(512, L, generator=g) * 0.2
    sc = torch.rand(512, generator=g) + 0.5
    sf = torch.randn(512, generator=g) * 0.1
    zz = z_r.reshape(-1, L)
    ref = F.relu((zz @ wz.t()) * sc + sf)
    got = ops.latent_project(zz.to(d).contiguous(), wz.to(d), sc.to(d), sf.to(d), True)
    assert (got.cpu() - ref).abs().max() < 1e-4 * max(1.0, ref.abs().max().item())


def test_final_conv_and_resize(cuda_dev):
    from vaeunet_b200 import ops
    g = torch.Generator().manual_seed(6)
    x = _q(torch.randn(3, 64, 20, 24, generator=g))
    w = torch.randn(64, generator=g) * 0.2
    ref = F.conv2d(x, w.view(1, 64, 1, 1), torch.tensor([0.3]))[:, 0]
    got = ops.final_conv1(_to_act(ops, x, cuda_dev), w.to(cuda_dev), 0.3)
    assert (got.cpu() - ref).abs().max() < 1e-4 * max(1.0, ref.abs().max().item())
    up = F.interpolate(ref[:, None], size=(40, 48), mode="bilinear", align_corners=True)[:, 0]
    a = ops.resize_logits(got, 40, 48, ops.SIG_NONE).cpu()
    b = ops.resize_logits(got, 40, 48, ops.SIG_POST).cpu()
    c = ops.resize_logits(got, 40, 48, ops.SIG_PRE).cpu()
    assert (a - up).abs().max() < 1e-4
    assert (b - torch.sigmoid(up)).abs().max() < 1e-5
    pre = F.interpolate(torch.sigmoid(ref)[:, None], size=(40, 48), mode="bilinear", align_corners=True)[:, 0]
    assert (c - pre).abs().max() < 1e-5


@pytest.mark.parametrize("Ns", [1, 2, 10])
def test_uncertainty_reduce(cuda_dev, Ns):
    from vaeunet_b200 import ops
    from oracle import vaeunet_oracle as O
    g = torch.Generator().manual_seed(7)
    B, Hs, Ws, Ho, Wo = 2, 16, 20, 32, 40
    lg = torch.randn(B, Ns, Hs, Ws, generator=g) * 3
    up = F.interpolate(lg, size=(Ho, Wo), mode="bilinear", align_corners=True)
    segs = torch.sigmoid(up).permute(1, 0, 2, 3).unsqueeze(2)  # [N,B,1,H,W]
    ref = O.calculate_uncertainty_metrics(segs)
    res = ops.uncertainty_reduce(lg.to(cuda_dev), Ho, Wo, want_probs=True)
    assert (res["samples"].cpu() - torch.sigmoid(up)).abs().max() < 1e-5
    for k in ("mean", "std", "entropy", "mutual_info", "coeff_var"):
        a, b = res[k].cpu(), ref[k]
        if Ns == 1 and k in ("std", "coeff_var"):
            assert torch.isnan(a).all() and torch.isnan(b).all()
            continue
        tol = 1e-4 * max(1.0, b.abs().max().item()) if k != "coeff_var" else 1e-3 * b.abs().max().item()
        assert (a - b).abs().max() < tol, k
    # probability input at output size
    res2 = ops.uncertainty_reduce(torch.sigmoid(up).to(cuda_dev).contiguous(), Ho, Wo, input_is_prob=True)
    assert (res2["mean"].cpu() - ref["mean"]).abs().max() < 1e-5


def test_moments_merge_matches_direct(cuda_dev):
    from vaeunet_b200 import ops
    from oracle import vaeunet_oracle as O
    g = torch.Generator().manual_seed(8)
    probs = torch.rand(15, 1, 30, 34, generator=g)
    ref = O.calculate_uncertainty_metrics(probs.unsqueeze(2))
    shards = [2, 2, 2, 2, 2, 2, 2, 1]
    means, m2s, shs = [], [], []
    o = 0
    for c in shards:
        m, m2, sh = ops.moments_partial(probs[o:o + c].to(cuda_dev).contiguous())
        means.append(m); m2s.append(m2); shs.append(sh)
        o += c
    res = ops.moments_finalize(shards, torch.stack(means), torch.stack(m2s), torch.stack(shs))
    for k in ("mean", "std", "entropy", "mutual_info"):
        assert (res[k].cpu() - ref[k]).abs().max() < 1e-4, k


def test_blend(cuda_dev):
    from vaeunet_b200 import ops
    from oracle import vaeunet_oracle as O
    H, W, ps, ov = 150, 170, 64, 16
    nh, nw, info = O.patch_grid(H, W, ps, ov)
    g = torch.Generator().manual_seed(9)
    acc_r = torch.zeros(H, W); ws_r = torch.zeros(H, W)
    acc = torch.zeros(H, W, device=cuda_dev); ws = torch.zeros(H, W, device=cuda_dev)
    for idx, (sh, eh, sw, ew) in enumerate(info):
        t = torch.rand(eh - sh, ew - sw, generator=g)
        i, j = divmod(idx, nw)
        wt = O.blend_weight(eh - sh, ew - sw, i, j, nh, nw, ov)
        acc_r[sh:eh, sw:ew] += t * wt; ws_r[sh:eh, sw:ew] += wt
        ops.blend_accumulate(t.to(cuda_dev), sh, sw, i, j, nh, nw, ov, acc, ws)
    out = ops.blend_finalize(acc, ws).cpu()
    assert (ws.cpu() - ws_r).abs().max() < 1e-6
    assert (out - acc_r / (ws_r + 1e-8)).abs().max() < 1e-5


def test_losses(cuda_dev):
    from vaeunet_b200 import ops
    from oracle import vaeunet_oracle as O
    g = torch.Generator().manual_seed(10)
    x = (torch.randn(4, 1, 64, 64, generator=g) * 3).requires_grad_(True)
    t = (torch.rand(4, 1, 64, 64, generator=g) > 0.9).float()
    d = cuda_dev
    s = ops.seg_loss_sums(x.detach().to(d), t.to(d), False, 0.9, 2.0).cpu()
    n = x.numel()
    smooth = 1.0
    dice = 1 - (2 * s[2] + smooth) / (max(s[0].item(), 0.5) + max(s[1].item(), 0.5) + smooth)
    comb = 0.5 * s[3] / n + 0.5 * dice
    ref_c = O.combined_loss(x, t)
    assert abs(comb.item() - ref_c.item()) < 1e-5
    ma = 0.5 * dice + 0.5 * s[4] / n
    ref_m = O.ma_segmentation_loss(x, t)
    assert abs(ma.item() - ref_m.item()) < 1e-5
    gc, = torch.autograd.grad(ref_c, x, retain_graph=True)
    gm, = torch.autograd.grad(ref_m, x)
    sums_d = ops.seg_loss_sums(x.detach().to(d), t.to(d), False, 0.9, 2.0)
    dc = ops.seg_loss_bwd(x.detach().to(d), t.to(d), sums_d, 0.5, 0.5, 0.0).cpu()
    dm = ops.seg_loss_bwd(x.detach().to(d), t.to(d), sums_d, 0.0, 0.5, 0.5, 0.9, 2.0).cpu()
    assert (dc - gc).abs().max() < 1e-4 * gc.abs().max() + 1e-9
    assert (dm - gm).abs().max() < 1e-3 * gm.abs().max() + 1e-9
    # brier / nll on probabilities
    p = torch.sigmoid(x.detach())
    sp = ops.seg_loss_sums(p.to(d), t.to(d), True).cpu()
    assert abs(sp[6].item() / n - O.brier_score(p, t)) < 1e-6
    assert abs(sp[7].item() / n - O.negative_log_likelihood(p, t)) < 1e-5


def test_kl_free_bits(cuda_dev):
    from vaeunet_b200 import ops
    from oracle import vaeunet_oracle as O
    g = torch.Generator().manual_seed(11)
    mu = (torch.randn(16, 32, generator=g) * 0.5).requires_grad_(True)
    lv = (torch.randn(16, 32, generator=g) * 0.3).requires_grad_(True)
    ref = O.kl_with_free_bits(mu, lv, 1e-3)
    gm, gl = torch.autograd.grad(ref, [mu, lv])
    kl, dmu, dlv = ops.kl_free_bits(mu.detach().to(cuda_dev), lv.detach().to(cuda_dev), 1e-3, want_grad=True)
    assert abs(kl.item() - ref.item()) < 1e-5
    assert (dmu.cpu() - gm).abs().max() < 1e-6 and (dlv.cpu() - gl).abs().max() < 1e-6


def test_ece_and_confusion(cuda_dev):
    from vaeunet_b200 import ops
    from oracle import vaeunet_oracle as O
    g = torch.Generator().manual_seed(12)
    p = torch.rand(3, 50, 60, generator=g)
    p[0, 0, :5] = 1.0  # p == 1.0 falls in no bin (uncertainty_metrics.py:122)
    p[0, 1, :5] = 0.0
    gt = (torch.rand(3, 50, 60, generator=g) > 0.7).float()
    ece_r, accs, confs, counts = O.expected_calibration_error(p, gt)
    h = ops.ece_hist(p.to(cuda_dev), gt.to(cuda_dev), 10).cpu()
    assert torch.equal(h[0].float(), torch.tensor(counts))
    cnt = h[0].clamp(min=1)
    ece = ((h[0] * (h[1] / cnt - h[2] / cnt).abs()).sum() / h[0].sum()).item()
    assert abs(ece - ece_r) < 1e-6
    c = ops.confusion_counts(p.to(cuda_dev), gt.to(cuda_dev), 0.5).cpu()
    pb = p > 0.5
    tb = gt > 0.5
    assert c.tolist() == [float((pb & tb).sum()), float((pb & ~tb).sum()), float((~pb & ~tb).sum()), float((~pb & tb).sum())]


def test_staged_host_pipeline(cuda_dev):
    """pipeline.StagedInput / StagedOutput: double-buffered copies on side streams deliver the right data in order."""
    from vaeunet_b200.pipeline import StagedInput, StagedOutput
    inp, outp = StagedInput(cuda_dev, 2), StagedOutput(cuda_dev, 2)
    host = [torch.full((4, 3, 64, 64), float(i)).pin_memory() for i in range(7)]
    got = []
    for i, h in enumerate(host):
        x, = inp.push(h)
        if i + 1 < len(host):
            inp.prefetch(host[i + 1])
        y = x * 2 + 1                                   # stand-in for the device step (current stream)
        for _ in range(20):
            y = y + 0
        inp.release()
        outp.push({"y": y}, on_ready=lambda d: got.append(d["y"].clone()))
    outp.flush()
    assert len(got) == len(host)
    for i, g in enumerate(got):
        assert torch.equal(g, torch.full((4, 3, 64, 64), 2.0 * i + 1)), i
