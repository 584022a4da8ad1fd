"""GPU parity of the training kernels (through the C ABI) against torch CPU autograd on the same
16-bit-rounded operands: conv weight gradient (MN-major tcgen05 GEMM over pixels) and data gradient."""
import zlib
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

CASES = [
    # name, N, H, W, Cin, Cout, k, stride
    ("wg_3x3_64_64", 2, 32, 32, 64, 64, 3, 1),
    ("wg_3x3_224_64_wide", 1, 24, 130, 224, 64, 3, 1),
    ("wg_3x3_128_256", 3, 16, 16, 128, 256, 3, 1),
    ("wg_1x1_512_128", 2, 16, 16, 512, 128, 1, 1),
    ("wg_3x3_s2_64_128", 2, 32, 32, 64, 128, 3, 2),
    ("wg_1x1_s2_64_128", 2, 32, 32, 64, 128, 1, 2),
    ("wg_3x3_32_32_odd", 2, 11, 13, 32, 32, 3, 1),
    ("wg_3x3_ksplit", 4, 128, 128, 64, 64, 3, 1),
    # Cin % 256 == 0: the 256-column N tile of the per-tap kernel
    ("wg_3x3_256_256", 4, 32, 32, 256, 256, 3, 1),
    ("wg_3x3_s2_256_512", 3, 32, 32, 256, 512, 3, 2),
    ("wg_3x3_512_512_small", 5, 16, 16, 512, 512, 3, 1),
]


def _setup(case, dev, dt=torch.bfloat16):
    name, N, H, W, Cin, Cout, k, stride = case
    g = torch.Generator().manual_seed(zlib.crc32(name.encode()))
    pad = k // 2
    x = torch.randn(N, Cin, H, W, generator=g).to(dt).float()
    w = (torch.randn(Cout, Cin, k, k, generator=g) / (Cin * k * k) ** 0.5).to(dt).float()
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    dy = torch.randn(N, Cout, Ho, Wo, generator=g).to(dt).float()
    return x, w, dy, pad


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_conv_wgrad(case, cuda_dev, monkeypatch):
    if case[4] % 256 == 0:
        monkeypatch.setenv("VU_WGRAD_BN256", "1")     # the 256-column N tile (off by default; read per call)
    from vaeunet_b200 import ops
    name, N, H, W, Cin, Cout, k, stride = case
    x, w, dy, pad = _setup(case, cuda_dev)
    wr = w.clone().requires_grad_(True)
    F.conv2d(x, wr, None, stride, pad).backward(dy)
    ref = wr.grad
    xa = ops.Act(x.permute(0, 2, 3, 1).contiguous().bfloat16().to(cuda_dev))
    dya = ops.Act(dy.permute(0, 2, 3, 1).contiguous().bfloat16().to(cuda_dev))
    got = ops.unpack_wgrad(ops.conv2d_wgrad(xa, dya, k, stride, pad), k).cpu()
    rel = ((got - ref).norm() / ref.norm()).item()
    assert rel < 2e-4, f"{name}: rel {rel}"           # fp32 accumulation of exact bf16 products: only summation order differs
    assert (got - ref).abs().max().item() <= 1e-3 * ref.abs().max().item() + 1e-4


def test_conv_wgrad_row_kernel(cuda_dev, monkeypatch):
    """The filter-row kernel (3 taps per work item from one halo'd X tile) forced on for small shapes, incl. ragged tiles."""
    from vaeunet_b200 import ops
    monkeypatch.setenv("VU_WGRAD_ROW3", "1")
    for case in [("r3_64_64", 2, 32, 32, 64, 64, 3, 1), ("r3_224_64", 1, 24, 130, 224, 64, 3, 1), ("r3_128_256", 3, 16, 16, 128, 256, 3, 1),
                 ("r3_ksplit", 4, 128, 128, 64, 64, 3, 1), ("r3_ragged", 2, 20, 48, 32, 96, 3, 1),
                 # Cout <= 64: two filter rows stacked in M (the shifted dY rows as the upper half of the A tile)
                 ("r3_stack_cout32", 2, 40, 64, 64, 32, 3, 1), ("r3_stack_nt2", 4, 32, 16, 96, 64, 3, 1),
                 ("r3_stack_ragged", 1, 37, 150, 160, 64, 3, 1)]:
        name, N, H, W, Cin, Cout, k, stride = case
        x, w, dy, pad = _setup(case, cuda_dev)
        wr = w.clone().requires_grad_(True)
        F.conv2d(x, wr, None, stride, pad).backward(dy)
        xa = ops.Act(x.permute(0, 2, 3, 1).contiguous().bfloat16().to(cuda_dev))
        dya = ops.Act(dy.permute(0, 2, 3, 1).contiguous().bfloat16().to(cuda_dev))
        got = ops.unpack_wgrad(ops.conv2d_wgrad(xa, dya, k, stride, pad), k).cpu()
        assert ((got - wr.grad).norm() / wr.grad.norm()).item() < 2e-4, name


def test_conv_epilogue_bn_statistics(cuda_dev):
    """conv2d(stat_sums=): the per-channel sum / sum of squares accumulated in the conv epilogue equal the statistics
    of the STORED (16-bit) output, on the tap-reuse kernel, the implicit GEMM (3x3 s1/s2, 1x1, bias, ragged pixel and
    channel tiles) alike. Checked against fp64 sums of the output read back."""
    from vaeunet_b200 import ops
    g = torch.Generator().manual_seed(5)
    for name, N, H, W, Cin, Cout, k, stride, bias in [("w64_c64", 2, 64, 64, 64, 64, 3, 1, False), ("w160_c32", 1, 96, 160, 96, 32, 3, 1, False),
                                                      # the smem tap-reuse kernel in its statistics flavour (W >= 96, Cout 64 / 128)
                                                      ("halo64_w128", 2, 10, 128, 64, 64, 3, 1, False), ("halo224_w256", 1, 7, 256, 224, 64, 3, 1, False),
                                                      ("halo128_w128", 3, 12, 128, 128, 128, 3, 1, False), ("halo64_odd", 2, 35, 175, 96, 64, 3, 1, False),
                                                      ("ig128", 2, 32, 32, 64, 128, 3, 1, False), ("ig_s2", 2, 33, 31, 64, 128, 3, 2, False),
                                                      ("ig_1x1_bias", 3, 20, 28, 128, 64, 1, 1, True), ("ig_320", 1, 16, 16, 64, 320, 3, 1, False),
                                                      ("ig_512", 4, 8, 8, 256, 512, 3, 1, False)]:
        for dt, tdt in ((ops.BF16, torch.bfloat16), (ops.F16, torch.float16)):
            x = torch.randn(N, H, W, Cin, generator=g).to(tdt).to(cuda_dev)
            w = (torch.randn(Cout, Cin, k, k, generator=g) / (Cin * k * k) ** 0.5).to(cuda_dev)
            b = torch.randn(Cout, generator=g).to(cuda_dev) if bias else None
            wp = ops.pack_conv_weight(w, dt)
            sums = torch.zeros(2 * Cout, dtype=torch.float64, device=cuda_dev)
            y = ops.conv2d(ops.Act(x), wp, k, stride, k // 2, None, b, False, stat_sums=sums)
            y0 = ops.conv2d(ops.Act(x), wp, k, stride, k // 2, None, b, False)
            assert torch.equal(y.t, y0.t), name                       # the statistics do not disturb the output
            yd = y.t.double().reshape(-1, Cout)
            ref = torch.cat([yd.sum(0), (yd * yd).sum(0)])
            scale = torch.cat([yd.abs().sum(0), (yd * yd).sum(0)]) + 1e-9
            err = ((sums - ref).abs() / scale).max().item()
            assert err < 2e-6, f"{name} {tdt}: {err}"                  # fp32 partials per CTA, fp64 across CTAs


@pytest.mark.parametrize("case", [c for c in CASES if c[2] % 2 == 0 or c[7] == 1], ids=lambda c: c[0].replace("wg", "dg"))
def test_conv_dgrad(case, cuda_dev):
    from vaeunet_b200 import ops
    name, N, H, W, Cin, Cout, k, stride = case
    x, w, dy, pad = _setup(case, cuda_dev)
    xr = x.clone().requires_grad_(True)
    F.conv2d(xr, w, None, stride, pad).backward(dy)
    ref = xr.grad
    dya = ops.Act(dy.permute(0, 2, 3, 1).contiguous().bfloat16().to(cuda_dev))
    banks = ops.pack_dgrad_weights(w.to(cuda_dev), stride, ops.BF16)
    got = ops.conv2d_dgrad(dya, banks, k, stride, H, W)
    g = got.t.float().cpu().permute(0, 3, 1, 2)
    rel = ((g - ref).norm() / ref.norm()).item()
    assert rel < 6e-3, f"{name}: rel {rel}"           # output rounded to bf16
    if stride == 1:                                    # accumulate form: dX += dgrad
        got2 = ops.conv2d_dgrad(dya, banks, k, stride, H, W, out=got, accumulate=True)
        g2 = got2.t.float().cpu().permute(0, 3, 1, 2)
        assert ((g2 - 2 * ref).norm() / (2 * ref).norm()).item() < 8e-3


# ---------------------------------------------------------------------------------------------------------
# memory-bound training kernels vs torch CPU autograd (fp32 reference on the same bf16-rounded inputs)
# ---------------------------------------------------------------------------------------------------------
def _act(ops, t, dev):
    return ops.Act(t.permute(0, 2, 3, 1).contiguous().bfloat16().to(dev))


def _nchw(a):
    return a.t[..., a.coff:a.coff + a.C].float().cpu().permute(0, 3, 1, 2)


def _q(t):
    return t.bfloat16().float()


@pytest.mark.parametrize("C,relu,res", [(64, True, False), (224, True, True), (8, False, False), (512, True, True)])
def test_bn_train_fwd_bwd(cuda_dev, C, relu, res):
    from vaeunet_b200 import ops, train_ops as T
    g = torch.Generator().manual_seed(C)
    N, H, W = 3, 9, 7
    x = _q(torch.randn(N, C, H, W, generator=g) * 2 + 0.5)
    r = _q(torch.randn(N, C, H, W, generator=g)) if res else None
    gamma = torch.rand(C, generator=g) + 0.5
    beta = torch.randn(C, generator=g) * 0.2
    dy = _q(torch.randn(N, C, H, W, generator=g))
    xr = x.clone().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    rm, rv = torch.zeros(C), torch.ones(C)
    y = F.batch_norm(xr, rm, rv, gr, br, True, 0.1, 1e-5)
    if res:
        y = y + r
    if relu:
        y = F.relu(y)
    y.backward(dy)
    d = cuda_dev
    rm_g, rv_g = torch.zeros(C, device=d), torch.ones(C, device=d)
    xa = _act(ops, x, d)
    out, mean, invstd = T.bn_train_fwd(xa, gamma.to(d), beta.to(d), rm_g, rv_g, relu, _act(ops, r, d) if res else None)
    assert (_nchw(out) - y.detach()).abs().max() <= 1.2e-2 * y.detach().abs().max()
    assert (rm_g.cpu() - rm).abs().max() < 1e-5 and (rv_g.cpu() - rv).abs().max() < 1e-4
    dx, dgam, dbet, dz = T.bn_train_bwd(xa, _act(ops, dy, d), out if relu else None, mean, invstd, gamma.to(d), want_dz=res)
    # the GPU ReLU mask comes from the bf16-rounded output; compare away from the mask boundary through norms
    assert ((_nchw(dx) - xr.grad).norm() / xr.grad.norm()).item() < 1.5e-2
    assert ((dgam.cpu() - gr.grad).norm() / gr.grad.norm()).item() < 5e-3
    assert ((dbet.cpu() - br.grad).norm() / br.grad.norm()).item() < 5e-3
    if relu and not res:     # ReLU mask recomputed from x (no read of y): identical result
        dx2, dgam2, dbet2, _ = T.bn_train_bwd(xa, _act(ops, dy, d), None, mean, invstd, gamma.to(d), relu_beta=beta.to(d))
        assert (dx2.t.float() - dx.t.float()).abs().max() <= 1e-2 * dx.t.float().abs().max()      # fp64 atomics order
        assert torch.allclose(dgam2, dgam, rtol=1e-5, atol=1e-6) and torch.allclose(dbet2, dbet, rtol=1e-5, atol=1e-6)
    if res:
        mask = (y.detach() > 0).float() if relu else torch.ones_like(dy)
        assert ((_nchw(dz) - dy * mask).norm() / (dy * mask).norm()).item() < 5e-3


def test_maxpool_bilinear_bwd(cuda_dev):
    from vaeunet_b200 import ops, train_ops as T
    g = torch.Generator().manual_seed(9)
    d = cuda_dev
    x = _q(torch.randn(2, 16, 12, 14, generator=g))
    xr = x.clone().requires_grad_(True)
    y = F.max_pool2d(xr, 3, 2, 1)
    dy = _q(torch.randn(y.shape, generator=g))
    y.backward(dy)
    dx = T.maxpool_bwd(_act(ops, x, d), _act(ops, dy, d), 3, 2, 1)
    assert (_nchw(dx) - xr.grad).abs().max() <= 1e-2 * xr.grad.abs().max()
    # saved-argmax variant (training fast path), with ties: ReLU'd input has many all-zero windows
    # (64 channels: the k3 s2 p1 specialisation of the backward; 16: the generic kernel)
    for (Cc, hh, ww) in [(16, 13, 14), (64, 13, 14), (64, 32, 40)]:
        xz = _q(torch.relu(torch.randn(2, Cc, hh, ww, generator=g) - 0.8))
        xzr = xz.clone().requires_grad_(True)
        yz = F.max_pool2d(xzr, 3, 2, 1)
        dyz = _q(torch.randn(yz.shape, generator=g))
        yz.backward(dyz)
        y2, code = T.maxpool_idx_fwd(_act(ops, xz, d), 3, 2, 1)
        assert (_nchw(y2) - yz.detach()).abs().max() == 0
        dx2 = T.maxpool_idx_bwd(code, _act(ops, dyz, d), hh, ww, 3, 2, 1)
        assert (_nchw(dx2) - xzr.grad).abs().max() <= 1e-2 * xzr.grad.abs().max(), (Cc, hh, ww)
    for (hs, ws, ho, wo) in [(8, 8, 16, 16), (11, 9, 22, 17), (1, 1, 5, 6), (16, 16, 32, 32), (64, 64, 128, 128),
                             (5, 7, 17, 20), (32, 48, 63, 97)]:     # x2 (unrolled tap kernel), x3 and odd sizes
        for Cc in (24, 64, 256):                                     # generic tap kernel / the pixel-run kernel
            x = torch.randn(2, Cc, hs, ws, generator=g).requires_grad_(True)
            y = F.interpolate(x, size=(ho, wo), mode="bilinear", align_corners=True)
            dy = _q(torch.randn(y.shape, generator=g))
            y.backward(dy)
            dx = T.bilinear_ac_bwd(_act(ops, dy, d), hs, ws)
            assert ((_nchw(dx) - x.grad).norm() / x.grad.norm()).item() < 5e-3, (hs, ws, ho, wo, Cc)
        m = torch.randn(3, ho, wo, generator=g)
        xm = torch.randn(3, 1, hs, ws, generator=g).requires_grad_(True)
        F.interpolate(xm, size=(ho, wo), mode="bilinear", align_corners=True).backward(m.unsqueeze(1))
        got = T.resize_map_bwd(m.to(d), hs, ws).cpu()
        assert (got - xm.grad[:, 0]).abs().max() < 1e-5


def test_gate_pieces_and_vec_layers(cuda_dev):
    from vaeunet_b200 import ops, train_ops as T
    g = torch.Generator().manual_seed(10)
    d = cuda_dev
    N, C, H, W = 2, 64, 6, 5
    a, b = _q(torch.randn(N, C, H, W, generator=g)), _q(torch.randn(N, C, H, W, generator=g))
    m = torch.rand(N, H, W, generator=g)
    out = ops.new_act(N, H, W, C)
    T.scale_by_map(_act(ops, a, d), m.to(d), out)
    assert (_nchw(out) - a * m.unsqueeze(1)).abs().max() < 2e-2
    assert (T.channel_dot(_act(ops, a, d), _act(ops, b, d)).cpu() - (a * b).sum(1)).abs().max() < 1e-3
    w = torch.randn(C, generator=g)
    T.map_expand(m.to(d), w.to(d), _act(ops, a, d), out)
    ref = m.unsqueeze(1) * w.view(1, -1, 1, 1) * (a > 0)
    assert (_nchw(out) - ref).abs().max() <= 1e-2 * ref.abs().max()
    s = T.weighted_channel_sum(_act(ops, a, d), m.to(d), per_image=False).cpu()
    assert (s - (a * m.unsqueeze(1)).sum((0, 2, 3))).abs().max() < 1e-3
    s = T.weighted_channel_sum(_act(ops, a, d), None, per_image=True).cpu()
    assert (s - a.sum((2, 3))).abs().max() < 1e-3
    # psi branch: BN(1 channel) + sigmoid
    p_raw = torch.randn(N, H, W, generator=g) * 2 + 1
    pr = p_raw.clone().requires_grad_(True)
    gam = torch.tensor([1.3], requires_grad=True); bet = torch.tensor([-0.2], requires_grad=True)
    rm, rv = torch.zeros(1), torch.ones(1)
    psi_ref = torch.sigmoid(F.batch_norm(pr.unsqueeze(1), rm, rv, gam, bet, True, 0.1, 1e-5))[:, 0]
    dpsi = torch.randn(N, H, W, generator=g)
    psi_ref.backward(dpsi)
    rm_g, rv_g = torch.zeros(1, device=d), torch.ones(1, device=d)
    gam_d, bet_d = torch.tensor([1.3], device=d), torch.tensor([-0.2], device=d)
    psi, stat = T.map_bn_sigmoid_fwd(p_raw.to(d), gam_d, bet_d, rm_g, rv_g)
    assert (psi.cpu() - psi_ref.detach()).abs().max() < 1e-5 and (rv_g.cpu() - rv).abs().max() < 1e-5
    dx, dgam, dbet = T.map_bn_sigmoid_bwd(p_raw.to(d), dpsi.to(d), psi, stat, gam_d)
    assert (dx.cpu() - pr.grad).abs().max() < 1e-5
    assert abs(dgam.item() - gam.grad.item()) < 1e-4 and abs(dbet.item() - bet.grad.item()) < 1e-4
    # dense vector layer with train-mode BN over the rows (z_initial / z_proj on constant maps)
    M, L, Cc = 6, 32, 80
    z = torch.randn(M, L, generator=g).requires_grad_(True)
    wv = (torch.randn(Cc, L, generator=g) * 0.3).requires_grad_(True)
    bv = torch.randn(Cc, generator=g).requires_grad_(True)
    gam = (torch.rand(Cc, generator=g) + 0.5).requires_grad_(True); bet = torch.randn(Cc, generator=g).requires_grad_(True)
    rm, rv = torch.zeros(Cc), torch.ones(Cc)
    y_ref = F.relu(F.batch_norm(F.linear(z, wv, bv), rm, rv, gam, bet, True, 0.1, 1e-5))
    dy = torch.randn(M, Cc, generator=g)
    y_ref.backward(dy)
    rm_g, rv_g = torch.zeros(Cc, device=d), torch.ones(Cc, device=d)
    y, saved = T.vec_layer_fwd(z.detach().to(d), wv.detach().to(d), bv.detach().to(d), gam.detach().to(d), bet.detach().to(d),
                               rm_g, rv_g, True, float(M))
    assert (y.cpu() - y_ref.detach()).abs().max() < 1e-4 and (rv_g.cpu() - rv).abs().max() < 1e-4
    dz = torch.zeros(M, L, device=d)
    dw, db, dg, dbt = T.vec_layer_bwd(dy.to(d), saved, wv.detach().to(d), gam.detach().to(d), True, dz, True)
    assert (dz.cpu() - z.grad).abs().max() < 1e-4 and (dw.cpu() - wv.grad).abs().max() < 1e-4
    assert (dg.cpu() - gam.grad).abs().max() < 1e-4 and (dbt.cpu() - bet.grad).abs().max() < 1e-4
    assert db.abs().max().item() < 1e-4          # a bias in front of train-mode BN has zero gradient


def test_stem_wgrad_and_adamw(cuda_dev):
    from vaeunet_b200 import ops, train_ops as T
    g = torch.Generator().manual_seed(11)
    d = cuda_dev
    x = torch.rand(2, 3, 40, 36, generator=g)
    w = (torch.randn(64, 3, 7, 7, generator=g) * 0.1).requires_grad_(True)
    y = F.conv2d(x, w, None, 2, 3)
    dy = _q(torch.randn(y.shape, generator=g))
    y.backward(dy)
    raw = T.stem_conv_raw(x.to(d), w.detach().to(d), ops.BF16)
    assert (_nchw(raw) - y.detach()).abs().max() <= 1e-2 * y.detach().abs().max()
    dw = T.stem_wgrad(x.to(d), _act(ops, dy, d), 64).cpu()
    assert ((dw - w.grad).norm() / w.grad.norm()).item() < 1e-4
    # tensor-core route: im2col patches (147 -> 152 columns) x 1x1 bank, forward and weight gradient
    xq = _q(x)
    yq = F.conv2d(xq, _q(w.detach()), None, 2, 3)
    wq = _q(w.detach()).clone().requires_grad_(True)
    F.conv2d(xq, wq, None, 2, 3).backward(dy)
    xcol = T.im2col_small_cin(x.to(d), 7, 2, 3, ops.BF16)
    assert xcol.C == 152
    w16 = T.pack_rows_padded(w.detach().to(d), 152, ops.BF16)
    y_tc = ops.conv2d(xcol, w16, 1, 1, 0, None, None, False)
    assert ((_nchw(y_tc) - yq).norm() / yq.norm()).item() < 4e-3
    dw_tc = ops.conv2d_wgrad(xcol, _act(ops, dy, d), 1, 1, 0).cpu().view(64, 152)
    assert dw_tc[:, 147:].abs().max().item() == 0
    assert ((dw_tc[:, :147].reshape(64, 3, 7, 7) - wq.grad).norm() / wq.grad.norm()).item() < 2e-4
    # AdamW with gradient clipping, two steps
    n = 1000
    p0 = torch.randn(n, generator=g); g1 = torch.randn(n, generator=g) * 3
    pt = p0.clone().requires_grad_(True)
    opt = torch.optim.AdamW([pt], lr=1e-3, weight_decay=1e-2)
    p, m, v = p0.clone().to(d), torch.zeros(n, device=d), torch.zeros(n, device=d)
    ss = torch.zeros(1, dtype=torch.float64, device=d)
    for step in (1, 2):
        gi = g1 * step
        pt.grad = gi.clone()
        torch.nn.utils.clip_grad_norm_([pt], 1.0)
        opt.step()
        T.sumsq(gi.to(d), ss, True)
        T.adamw_step(p, gi.to(d), m, v, 1e-3, (0.9, 0.999), 1e-8, 1e-2, step, ss, 1.0)
    assert (p.cpu() - pt.detach()).abs().max() < 1e-5


def test_batched_weight_packing(cuda_dev):
    """vu_pack_weights_batched (one launch per step for all 50 filter banks, forward + data-gradient layouts) is bit-identical
    to the single-tensor packers vu_pack_conv_weight / vu_pack_dgrad_weight for every convolution of the model."""
    from vaeunet_b200 import ops, testing
    from vaeunet_b200.training import TrainStep
    from vaeunet_b200.unet import UNetResNet
    sd, kw = testing.resnet_fixture()
    model = UNetResNet(**kw)
    model.load_state_dict(sd)
    model = model.to(cuda_dev).train()
    for dt in (ops.BF16, ops.F16):
        ts = TrainStep(model, dtype=dt, loss_scale=1.0)
        ts.repack()
        torch.cuda.synchronize()
        strides = set()
        for name, c in ts.convs.items():
            w = ts.P[c.name]
            assert torch.equal(c.w16, ops.pack_conv_weight(w, dt)), name
            ref = ops.pack_dgrad_weights(w, c.stride, dt)
            got = c.banks if isinstance(c.banks, (list, tuple)) else [c.banks]
            ref = ref if isinstance(ref, (list, tuple)) else [ref]
            for a, b in zip(got, ref):
                assert torch.equal(a.reshape(-1), b.reshape(-1)), name
            strides.add((c.k, c.stride))
        assert {(3, 1), (3, 2), (1, 1), (1, 2)} <= strides
