"""GPU parity of the tcgen05 implicit-GEMM conv (through the C ABI) against torch CPU fp32
F.conv2d on the same 16-bit-rounded operands (the oracle's primitive)."""
import zlib
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

CASES = [
    # name, N, H, W, Cin, Cout, k, stride, relu, residual, x_extra(coff,ctot+), y_extra, dtype
    ("3x3_64_64", 2, 32, 32, 64, 64, 3, 1, True, False, (0, 0), (0, 0), "bf16"),
    ("3x3_partialK_224", 1, 40, 40, 224, 64, 3, 1, True, False, (0, 0), (0, 0), "bf16"),
    ("3x3_800_512_res", 3, 16, 16, 800, 512, 3, 1, True, True, (0, 0), (0, 0), "bf16"),
    ("1x1_512_128", 2, 32, 32, 512, 128, 1, 1, False, False, (0, 0), (0, 0), "bf16"),
    ("3x3_s2_64_128", 2, 32, 32, 64, 128, 3, 2, True, False, (0, 0), (0, 0), "bf16"),
    ("3x3_s2_odd", 2, 35, 29, 64, 128, 3, 2, True, False, (0, 0), (0, 0), "bf16"),
    ("1x1_s2_ds", 2, 32, 32, 64, 128, 1, 2, False, False, (0, 0), (0, 0), "bf16"),
    ("3x3_slices", 2, 24, 24, 96, 64, 3, 1, True, False, (32, 160), (64, 192), "bf16"),
    ("3x3_tiny_maps", 5, 4, 4, 128, 256, 3, 1, True, True, (0, 0), (0, 0), "bf16"),
    ("3x3_cin32", 2, 16, 16, 32, 32, 3, 1, False, False, (0, 0), (0, 0), "bf16"),
    ("3x3_fp16", 2, 32, 32, 128, 128, 3, 1, True, False, (0, 0), (0, 0), "f16"),
    ("3x3_22x22", 2, 22, 22, 256, 256, 3, 1, True, False, (0, 0), (0, 0), "bf16"),
    ("3x3_1024", 1, 32, 32, 1024, 1024, 3, 1, True, False, (0, 0), (0, 0), "bf16"),
    ("3x3_big_persistent", 4, 128, 128, 64, 64, 3, 1, True, False, (0, 0), (0, 0), "bf16"),
    # smem tap-reuse kernel (vu_conv_halo.cu): Cout 64/128, W >= 96
    ("halo_64_ht4", 2, 64, 128, 64, 64, 3, 1, True, False, (0, 0), (0, 0), "f16"),
    ("halo_64_ht2", 3, 6, 128, 64, 64, 3, 1, False, False, (0, 0), (0, 0), "bf16"),
    ("halo_partialK_224_w256", 1, 40, 256, 224, 64, 3, 1, True, False, (0, 0), (0, 0), "f16"),
    ("halo_128_res", 2, 32, 128, 128, 128, 3, 1, True, True, (0, 0), (0, 0), "bf16"),
    ("halo_odd_slices", 2, 35, 175, 96, 64, 3, 1, True, True, (32, 160), (64, 192), "f16"),
    ("halo_352_128_w128", 1, 128, 128, 352, 128, 3, 1, True, False, (0, 0), (0, 0), "f16"),
    # CTA-pair kernel (256-pixel tiles over two SMs): odd number of 128-pixel tiles (the peer's last tile is empty), many
    # tiles per pair, partial last filter tile (Cout 320 with 256-column tiles), residual + channel slices
    ("pair_odd_tiles", 3, 16, 16, 256, 256, 3, 1, True, True, (0, 0), (0, 0), "bf16"),
    ("pair_odd_tiles_7", 7, 8, 16, 128, 128, 3, 1, True, False, (0, 0), (0, 0), "f16"),
    ("pair_persistent", 16, 64, 64, 128, 128, 3, 1, True, False, (0, 0), (0, 0), "bf16"),
    ("pair_cout320", 2, 32, 32, 128, 320, 3, 1, True, False, (0, 0), (0, 0), "bf16"),
    ("pair_slices_res", 2, 32, 32, 192, 256, 3, 1, True, True, (64, 320), (128, 512), "f16"),
    ("pair_1x1_s2", 4, 32, 32, 128, 256, 1, 2, False, False, (0, 0), (0, 0), "bf16"),
    # flat padded tap-reuse kernel (vu_conv_flat.cu): W <= 64; 256 / 128 positions per CTA, several N tiles, partial last
    # N tile (Cout 672 = 5 x 128 + 32), partial last K block (Cin 672), odd tile counts, ragged heights, slices + residual
    ("flat_w64_128", 16, 64, 64, 128, 128, 3, 1, True, False, (0, 0), (0, 0), "f16"),
    ("flat_w32_256", 16, 32, 32, 256, 256, 3, 1, True, True, (0, 0), (0, 0), "f16"),
    ("flat_w16_512", 8, 16, 16, 512, 512, 3, 1, True, False, (0, 0), (0, 0), "bf16"),
    ("flat_cout672", 2, 64, 64, 256, 672, 3, 1, False, False, (0, 0), (0, 0), "f16"),
    ("flat_cin672", 2, 64, 64, 672, 256, 3, 1, True, False, (0, 0), (0, 0), "f16"),
    ("flat_ragged", 3, 37, 41, 96, 160, 3, 1, True, True, (32, 160), (64, 256), "f16"),
    ("flat_w8", 5, 9, 8, 64, 64, 3, 1, True, False, (0, 0), (0, 0), "bf16"),
    ("flat_one_image", 1, 64, 64, 128, 128, 3, 1, True, False, (0, 0), (0, 0), "f16"),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_conv_igemm(case, cuda_dev, monkeypatch):
    from vaeunet_b200 import ops
    name, N, H, W, Cin, Cout, k, stride, relu, use_res, xs, ys, dt = case
    if name.startswith("flat_"):
        monkeypatch.setenv("VU_FLAT", "2")       # every shape that fits (auto mode takes Cout <= 128 only)
    tdt = torch.bfloat16 if dt == "bf16" else torch.float16
    odt = ops.BF16 if dt == "bf16" else ops.F16
    g = torch.Generator().manual_seed(zlib.crc32(name.encode()))
    pad = 1 if k == 3 else 0
    x = torch.randn(N, Cin, H, W, generator=g)
    w = torch.randn(Cout, Cin, k, k, generator=g) * (1.0 / (Cin * k * k) ** 0.5)
    scale = torch.rand(Cout, generator=g) + 0.5
    bias = torch.randn(Cout, generator=g) * 0.1
    xq = x.to(tdt).float()
    wq = w.to(tdt).float()
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    res = torch.randn(N, Cout, Ho, Wo, generator=g).to(tdt).float() if use_res else None
    ref = F.conv2d(xq, wq, None, stride, pad) * scale.view(1, -1, 1, 1) + bias.view(1, -1, 1, 1)
    if use_res:
        ref = ref + res
    if relu:
        ref = F.relu(ref)

    dev = cuda_dev
    x_ctot = xs[1] if xs[1] else Cin
    xbuf = torch.full((N, H, W, x_ctot), 7.0, dtype=tdt, device=dev)
    xbuf[..., xs[0]:xs[0] + Cin] = xq.permute(0, 2, 3, 1).to(tdt).to(dev)
    xa = ops.Act(xbuf, xs[0], Cin)
    y_ctot = ys[1] if ys[1] else Cout
    ybuf = torch.full((N, Ho, Wo, y_ctot), -3.0, dtype=tdt, device=dev)
    ya = ops.Act(ybuf, ys[0], Cout)
    wp = ops.pack_conv_weight(w.to(dev), odt)
    ra = ops.Act(res.permute(0, 2, 3, 1).contiguous().to(tdt).to(dev)) if use_res else None
    ops.conv2d(xa, wp, k, stride, pad, scale.to(dev), bias.to(dev), relu, out=ya, residual=ra)
    torch.cuda.synchronize()
    got = ybuf[..., ys[0]:ys[0] + Cout].float().cpu().permute(0, 3, 1, 2)
    err = (got - ref).abs().max().item()
    tol = 2e-2 * ref.abs().max().item()
    assert err <= tol, f"{name}: max err {err} > {tol}"
    rel = ((got - ref).norm() / ref.norm()).item()
    assert rel < 6e-3, f"{name}: rel l2 err {rel}"
    if ys[1]:  # untouched channels of the concat buffer stay untouched
        assert (ybuf[..., :ys[0]].float() == -3.0).all() and (ybuf[..., ys[0] + Cout:].float() == -3.0).all()


def test_convT2x2_pixel_shuffle(cuda_dev):
    """ConvTranspose2d k2 s2 (unet/unet_parts.py:76) as four 1x1 GEMMs with out_up=2."""
    from vaeunet_b200 import ops
    g = torch.Generator().manual_seed(5)
    N, Cin, Cout, H, W = 2, 128, 64, 16, 16
    x = torch.randn(N, Cin, H, W, generator=g)
    w = torch.randn(Cin, Cout, 2, 2, generator=g) * 0.1
    b = torch.randn(Cout, generator=g) * 0.1
    xq, wq = x.bfloat16().float(), w.bfloat16().float()
    ref = F.conv_transpose2d(xq, wq, b, stride=2)
    dev = cuda_dev
    xa = ops.Act(xq.permute(0, 2, 3, 1).contiguous().bfloat16().to(dev))
    wp = ops.pack_convT2x2_weight(w.to(dev))
    out = ops.new_act(N, 2 * H, 2 * W, Cout)
    for ab in range(4):
        ops.conv2d(xa, wp[ab], 1, 1, 0, None, b.to(dev), False, out=out, out_up=2, out_dh=ab // 2, out_dw=ab % 2)
    torch.cuda.synchronize()
    got = out.t.float().cpu().permute(0, 3, 1, 2)
    assert (got - ref).abs().max().item() <= 2e-2 * ref.abs().max().item()


@pytest.mark.parametrize("Cout,store", [(64, False), (64, True), (128, True)])
def test_conv_fused_output_conv(cuda_dev, Cout, store):
    """vu_conv_args.fuse_out: conv3x3 + BN + ReLU with the 1x1 Cout->1 output conv (final_conv) in the epilogue."""
    from vaeunet_b200 import ops
    g = torch.Generator().manual_seed(77)
    N, H, W, Cin = 2, 10, 130, 64
    x = torch.randn(N, Cin, H, W, generator=g)
    w = torch.randn(Cout, Cin, 3, 3, generator=g) * (1.0 / (Cin * 9) ** 0.5)
    scale = torch.rand(Cout, generator=g) + 0.5
    bias = torch.randn(Cout, generator=g) * 0.1
    fw = torch.randn(Cout, generator=g) * 0.2
    fb = 0.3
    xq, wq = x.half().float(), w.half().float()
    act = F.relu(F.conv2d(xq, wq, None, 1, 1) * scale.view(1, -1, 1, 1) + bias.view(1, -1, 1, 1))
    ref = (act * fw.view(1, -1, 1, 1)).sum(1) + fb
    dev = cuda_dev
    xa = ops.Act(xq.permute(0, 2, 3, 1).contiguous().half().to(dev))
    wp = ops.pack_conv_weight(w.to(dev), ops.F16)
    logits = torch.full((N, H, W), -9.0, device=dev)
    y = ops.conv2d(xa, wp, 3, 1, 1, scale.to(dev), bias.to(dev), True, fuse_w=fw.to(dev), fuse_bias=fb,
                   fuse_out=logits, store=store)
    torch.cuda.synchronize()
    assert (logits.cpu() - ref).abs().max().item() <= 5e-3 * ref.abs().max().item()
    if store:
        got = y.t.float().cpu().permute(0, 3, 1, 2)
        assert (got - act).abs().max().item() <= 2e-3 * act.abs().max().item() + 1e-3
    else:
        assert y is None


def test_conv_fused_output_rejected_off_path(cuda_dev):
    from vaeunet_b200 import ops
    from vaeunet_b200._lib import VaeUnetError
    dev = cuda_dev
    xa = ops.Act(torch.zeros(1, 8, 8, 64, dtype=torch.float16, device=dev))      # too narrow for the tap-reuse kernel
    wp = ops.pack_conv_weight(torch.zeros(64, 64, 3, 3, device=dev), ops.F16)
    with pytest.raises(VaeUnetError):
        ops.conv2d(xa, wp, 3, 1, 1, None, None, True, fuse_w=torch.zeros(64, device=dev), fuse_out=torch.zeros(1, 8, 8, device=dev))


@pytest.mark.parametrize("N,H,W,Cx,Cz,Cout", [(3, 20, 24, 192, 32, 64),      # implicit-GEMM kernel
                                              (2, 9, 130, 192, 32, 64),       # tap-reuse kernel, ragged tile
                                              (2, 32, 128, 320, 32, 128),     # tap-reuse kernel, Cout 128
                                              (2, 16, 16, 768, 32, 512)])     # 256-column tiles
def test_conv_pos_bias_constant_channels(cuda_dev, N, H, W, Cx, Cz, Cout):
    """conv3x3 over [x | spatially constant z] == conv3x3 over x with the z term as a border-class bias (pos_bias)."""
    from vaeunet_b200 import ops
    g = torch.Generator().manual_seed(N * 1000 + W)
    q = lambda t: t.half().float()
    x = q(torch.randn(N, Cx, H, W, generator=g))
    z = torch.randn(N, Cz, generator=g)
    w = torch.randn(Cout, Cx + Cz, 3, 3, generator=g) / (9 * (Cx + Cz)) ** 0.5
    w[:, :Cx] = q(w[:, :Cx])
    scale = torch.rand(Cout, generator=g) + 0.5
    bias = torch.randn(Cout, generator=g) * 0.1
    full = torch.cat([x, z.view(N, Cz, 1, 1).expand(N, Cz, H, W)], 1)
    ref = F.relu(F.conv2d(full, w, None, 1, 1) * scale.view(1, -1, 1, 1) + bias.view(1, -1, 1, 1))
    d = cuda_dev
    xa = ops.Act(x.permute(0, 2, 3, 1).contiguous().half().to(d))
    wp = ops.pack_conv_weight(w[:, :Cx].contiguous().to(d), ops.F16)
    pb = ops.border_bias(w.to(d), Cx, z.to(d))
    pb2 = ops.border_bias_apply(ops.border_wsum(w.to(d), Cx, Cz), z.to(d))        # two-step form used by the engine
    assert (pb - pb2).abs().max().item() < 1e-5
    y = ops.conv2d(xa, wp, 3, 1, 1, scale.to(d), bias.to(d), True, pos_bias=pb)
    got = y.t.float().cpu().permute(0, 3, 1, 2)
    assert ((got - ref).norm() / ref.norm()).item() < 1e-3
    assert (got - ref).abs().max().item() < 2e-3 * ref.abs().max().item() + 1e-3
