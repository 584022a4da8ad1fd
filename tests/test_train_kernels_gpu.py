"""GPU parity of the training kernels (through the C ABI) against torch CPU autograd on the same
16-bit-rounded operands: conv weight gradient (MN-major tcgen05 GEMM over pixels) and data gradient."""
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
]


def _setup(case, dev, dt=torch.bfloat16):
    name, N, H, W, Cin, Cout, k, stride = case
    g = torch.Generator().manual_seed(abs(hash(name)) % (2 ** 31))
    pad = k // 2
    x = torch.randn(N, Cin, H, W, generator=g).to(dt).float()
    w = (torch.randn(Cout, Cin, k, k, generator=g) / (Cin * k * k) ** 0.5).to(dt).float()
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    dy = torch.randn(N, Cout, Ho, Wo, generator=g).to(dt).float()
    return x, w, dy, pad


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_conv_wgrad(case, cuda_dev):
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
