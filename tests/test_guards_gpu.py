"""Out-of-bounds and race checks of our own (compute-sanitizer is closed on this GPU pool, profiles/r2za_sanitize_memcheck.txt):

  * guard bands: every output lives INSIDE a larger buffer pre-filled with a sentinel - guard elements before and after the
    tensor and, for NHWC outputs, sentinel channels on both sides of the written channel slice. After the kernel every
    sentinel must be intact (an out-of-bounds or mis-strided store of a TMA/tcgen05 epilogue, a ragged-tile store, a split-K
    reduction past the end would hit one) and the result must equal the one computed into a plain tensor;
  * run-to-run identity: kernels whose warps share shared-memory partials (BatchNorm statistics in the conv epilogues - the
    place where round 2's tap-reuse race lived) are run several times on the same inputs and must agree bit for bit in the
    stored outputs and to 1e-12 in the fp64 statistics.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

GUARD = 4096          # sentinel elements before and after every output


def _guarded(shape, dtype, dev, sentinel):
    n = 1
    for s in shape:
        n *= s
    buf = torch.full((n + 2 * GUARD,), sentinel, dtype=dtype, device=dev)
    view = buf[GUARD:GUARD + n].view(*shape)

    def intact():
        ref = torch.full((GUARD,), sentinel, dtype=dtype, device=dev)
        return bool(torch.equal(buf[:GUARD], ref) and torch.equal(buf[GUARD + n:], ref))
    return view, intact


def _slice_out(ops, N, H, W, C, dev, left=8, right=16):
    """An Act that is the channel slice [left, left+C) of a guarded [N,H,W,left+C+right] fp16 buffer filled with 12345."""
    t, intact = _guarded((N, H, W, left + C + right), torch.float16, dev, 12345.0)
    act = ops.Act(t, left, C)

    def ok():
        side = torch.full((1,), 12345.0, dtype=torch.float16, device=dev)
        return intact() and bool((t[..., :left] == side).all()) and bool((t[..., left + C:] == side).all())
    return act, ok


CONV_SHAPES = [
    # N, H, W, Cin, Cout, k, stats     kernel it lands on
    (2, 20, 24, 64, 32, 3, False),     # conv_igemm_kernel<32>
    (3, 33, 31, 128, 128, 3, False),   # pair<128>, ragged tiles
    (2, 16, 16, 256, 256, 3, True),    # pair<256> + statistics
    (1, 40, 130, 64, 64, 3, False),    # tap-reuse HT=4, ragged width
    (2, 50, 128, 64, 64, 3, True),     # tap-reuse HT=2 + statistics (the shape class of the round-2 race)
    (1, 37, 100, 128, 128, 3, True),   # tap-reuse<128> + statistics, ragged
    (2, 24, 24, 96, 64, 1, False),     # 1x1
    (16, 64, 64, 128, 128, 3, True),   # flat padded tap-reuse (W <= 64), 256 positions per CTA + statistics
    (16, 32, 32, 256, 256, 3, True),   # flat, two N tiles + statistics
    (3, 21, 30, 64, 96, 3, False),     # flat, ragged, partial N tile
]


@pytest.mark.parametrize("N,H,W,Cin,Cout,k,stats", CONV_SHAPES)
def test_conv_stays_inside_its_slice(cuda_dev, monkeypatch, N, H, W, Cin, Cout, k, stats):
    from vaeunet_b200 import ops
    if (N, H) in ((16, 64), (16, 32), (3, 21)):
        monkeypatch.setenv("VU_FLAT", "2")       # the flat padded tap-reuse kernel for every narrow shape (the other rows keep theirs)
    g = torch.Generator(device=cuda_dev).manual_seed(N * 1000 + H)
    x = ops.Act(torch.randn(N, H, W, Cin, device=cuda_dev, generator=g).half())
    w = ops.pack_conv_weight(torch.randn(Cout, Cin, k, k, device=cuda_dev, generator=g) * (Cin * k * k) ** -0.5, ops.F16)
    sc = torch.rand(Cout, device=cuda_dev, generator=g) + 0.5
    bi = torch.randn(Cout, device=cuda_dev, generator=g) * 0.1
    kw = dict(scale=None, bias=None, relu=False) if stats else dict(scale=sc, bias=bi, relu=True)
    s_ref = torch.zeros(2 * Cout, dtype=torch.float64, device=cuda_dev) if stats else None
    ref = ops.conv2d(x, w, k, 1, k // 2, kw["scale"], kw["bias"], kw["relu"], stat_sums=s_ref)
    out, ok = _slice_out(ops, N, H, W, Cout, cuda_dev)
    s_buf, s_ok = _guarded((2 * Cout,), torch.float64, cuda_dev, 7.0)
    if stats:
        s_buf.zero_()
    ops.conv2d(x, w, k, 1, k // 2, kw["scale"], kw["bias"], kw["relu"], out=out, stat_sums=s_buf if stats else None)
    torch.cuda.synchronize()
    assert ok(), "conv epilogue stored outside its channel slice / tensor"
    assert torch.equal(out.t[..., out.coff:out.coff + Cout], ref.t)
    if stats:
        assert s_ok(), "statistics flush outside [2*Cout]"
        assert torch.allclose(s_buf, s_ref, rtol=1e-12, atol=1e-9)


@pytest.mark.parametrize("N,H,W,Cin,Cout,k,stride,row3", [
    (2, 32, 32, 64, 64, 3, 1, True), (4, 128, 128, 64, 64, 3, 1, True),        # filter-row kernel, two rows stacked in M
    (1, 37, 150, 160, 64, 3, 1, True), (2, 64, 128, 96, 128, 3, 1, True),      # stacked + ragged / plain filter-row
    (3, 16, 16, 128, 256, 3, 1, False), (2, 32, 32, 64, 128, 3, 2, False),     # per-tap kernel, stride 2
    (2, 16, 16, 512, 128, 1, 1, False)])
def test_wgrad_stays_inside_its_buffer(cuda_dev, monkeypatch, N, H, W, Cin, Cout, k, stride, row3):
    from vaeunet_b200 import ops
    if row3:
        monkeypatch.setenv("VU_WGRAD_ROW3", "1")
    g = torch.Generator(device=cuda_dev).manual_seed(H + Cin)
    pad = k // 2
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    x = ops.Act(torch.randn(N, H, W, Cin, device=cuda_dev, generator=g).half())
    dy = ops.Act(torch.randn(N, Ho, Wo, Cout, device=cuda_dev, generator=g).half())
    ref = ops.conv2d_wgrad(x, dy, k, stride, pad)
    buf, ok = _guarded((Cout, k * k, Cin), torch.float32, cuda_dev, 3.0)
    buf.zero_()
    ops.conv2d_wgrad(x, dy, k, stride, pad, out=buf)
    torch.cuda.synchronize()
    assert ok(), "split-K reduction outside the packed gradient"
    # split-K sums are fp32 atomics: equal up to their order
    assert ((buf - ref).norm() / ref.norm()).item() < 1e-5


def test_concat_and_first_conv_stay_inside(cuda_dev):
    from vaeunet_b200 import ops
    g = torch.Generator(device=cuda_dev).manual_seed(5)
    r = lambda *s: ops.Act(torch.randn(*s, device=cuda_dev, generator=g).half())        # noqa: E731
    for (D, rep, hs, ws, H, W, Cx, Cs, F_) in [(6, 3, 9, 11, 18, 21, 128, 64, 32), (4, 2, 5, 6, 10, 12, 256, 64, 64),
                                               (4, 2, 3, 3, 6, 7, 512, 128, 128)]:
        x, g1 = r(D, hs, ws, Cx), r(D, hs, ws, F_)
        x1, skip = r(D // rep, H, W, F_), r(D // rep, H, W, Cs)
        pw = torch.randn(F_, device=cuda_dev, generator=g) * 0.2
        ref = ops.decoder_concat(x, g1, x1, skip, pw, 0.8, 0.05, None, ops.new_act(D, H, W, Cx + Cs, ops.F16, cuda_dev))
        t, ok = _guarded((D, H, W, Cx + Cs), torch.float16, cuda_dev, 12345.0)
        ops.decoder_concat(x, g1, x1, skip, pw, 0.8, 0.05, None, ops.Act(t))
        torch.cuda.synchronize()
        assert ok() and torch.equal(t, ref.t)
    for H, W in [(64, 64), (50, 70), (37, 129)]:
        img = torch.rand(2, 3, H, W, device=cuda_dev, generator=g)
        w = torch.randn(64, 3, 7, 7, device=cuda_dev, generator=g) * 0.1
        s = torch.rand(64, device=cuda_dev, generator=g) + 0.5
        b = torch.randn(64, device=cuda_dev, generator=g) * 0.1
        ref = ops.stem_conv7x7s2(img, w, s, b, ops.F16)
        out, ok = _slice_out(ops, 2, ref.H, ref.W, 64, cuda_dev)
        ops.stem_conv7x7s2(img, w, s, b, ops.F16, out=out)
        torch.cuda.synchronize()
        assert ok() and torch.equal(out.t[..., out.coff:out.coff + 64], ref.t)


@pytest.mark.parametrize("N,H,W,Cin,Cout", [(4, 128, 128, 64, 64), (2, 96, 256, 160, 64), (2, 128, 128, 128, 128), (8, 32, 32, 256, 256),
                                            (16, 64, 64, 128, 128), (16, 16, 16, 512, 512)])
def test_training_epilogue_is_run_to_run_identical(cuda_dev, monkeypatch, N, H, W, Cin, Cout):
    if N == 16:
        monkeypatch.setenv("VU_FLAT", "2")
    """The statistics partials are per warp; a shared slice raced in round 2 (wrong sums in ~1 of 3 runs at W >= 96)."""
    from vaeunet_b200 import ops
    g = torch.Generator(device=cuda_dev).manual_seed(W + Cin)
    x = ops.Act(torch.randn(N, H, W, Cin, device=cuda_dev, generator=g).half())
    w = ops.pack_conv_weight(torch.randn(Cout, Cin, 3, 3, device=cuda_dev, generator=g) * (Cin * 9) ** -0.5, ops.F16)
    outs, sums = [], []
    for _ in range(6):
        s = torch.zeros(2 * Cout, dtype=torch.float64, device=cuda_dev)
        outs.append(ops.conv2d(x, w, 3, 1, 1, None, None, False, stat_sums=s).t.clone())
        sums.append(s)
    torch.cuda.synchronize()
    y = outs[0].float()
    direct = torch.cat([y.sum((0, 1, 2)).double(), (y.double() ** 2).sum((0, 1, 2))])
    for o, s in zip(outs, sums):
        assert torch.equal(o, outs[0])
        assert torch.allclose(s, sums[0], rtol=1e-12, atol=1e-9)
        assert torch.allclose(s, direct, rtol=1e-5, atol=1e-3)
