"""torch-tensor wrappers over the training entry points of the C ABI (include/vaeunet_b200.h, "training step").
Same conventions as ops.py: `Act` = channel slice of an NHWC 16-bit buffer; fp32 tensors are contiguous CUDA."""
from __future__ import annotations

from typing import Optional

import torch

from . import ops
from ._lib import check, load, require_device
from .ops import Act, _count, _f32, _ptr, _stream

BN_EPS, BN_MOMENTUM = 1e-5, 0.1


def _npix(a: Act) -> int:
    return a.N * a.H * a.W


def _sums(dev, n=2048) -> torch.Tensor:
    return torch.empty(n, dtype=torch.float64, device=dev)


def bn_train_fwd(x: Act, gamma, beta, running_mean, running_var, relu: bool, residual: Optional[Act] = None,
                 out: Optional[Act] = None, update_running: bool = True, sums: Optional[torch.Tensor] = None):
    """Train-mode BatchNorm2d (+residual, +ReLU). -> (y, mean, invstd).
    sums: fp64 [2C] per-channel sum / sum of squares of x already accumulated by the producing conv
    (ops.conv2d(stat_sums=)); the statistics pass over x is skipped."""
    require_device()
    Cc, dev = x.C, x.t.device
    if sums is None:
        sums = _sums(dev, 2 * Cc)
        check(load().vu_bn_stats(x.t.data_ptr(), _npix(x), Cc, x.ctot, x.coff, sums.data_ptr(), x.dt, _stream()), "vu_bn_stats")
        _count(2)
    else:
        assert sums.dtype == torch.float64 and sums.numel() == 2 * Cc
    mean = torch.empty(Cc, dtype=torch.float32, device=dev)
    invstd = torch.empty(Cc, dtype=torch.float32, device=dev)
    if out is None:
        out = ops.new_act(x.N, x.H, x.W, Cc, x.dt, dev)
    r = residual
    # finalize (mean / invstd / running statistics) happens inside the apply kernel: one launch
    check(load().vu_bn_apply_sums(x.t.data_ptr(), _npix(x), Cc, x.ctot, x.coff, sums.data_ptr(), BN_EPS, BN_MOMENTUM,
                                  _ptr(running_mean) if update_running else None,
                                  _ptr(running_var) if update_running else None, mean.data_ptr(), invstd.data_ptr(),
                                  _f32(gamma).data_ptr(), _f32(beta).data_ptr(), r.t.data_ptr() if r else None,
                                  r.ctot if r else 8, r.coff if r else 0, int(relu), out.t.data_ptr(), out.ctot, out.coff, x.dt,
                                  _stream()), "vu_bn_apply_sums")
    _count(1)
    return out, mean, invstd


def bn_update_running(mean: torch.Tensor, invstd: torch.Tensor, count: float, running_mean: torch.Tensor,
                      running_var: torch.Tensor) -> None:
    """running statistics of nn.BatchNorm2d (momentum 0.1, unbiased variance) from saved batch statistics."""
    require_device()
    check(load().vu_bn_update_running(_f32(mean).data_ptr(), _f32(invstd).data_ptr(), float(count), mean.numel(), BN_EPS,
                                      BN_MOMENTUM, _f32(running_mean).data_ptr(), _f32(running_var).data_ptr(), _stream()),
          "vu_bn_update_running")
    _count()


def bn_train_bwd(x_raw: Act, dy: Act, y_relu: Optional[Act], mean, invstd, gamma, want_dz: bool = False,
                 dx: Optional[Act] = None, dgamma: Optional[torch.Tensor] = None, dbeta: Optional[torch.Tensor] = None,
                 relu_beta: Optional[torch.Tensor] = None):
    """-> (dx_raw, dgamma, dbeta, dz or None). y_relu: the post-ReLU output (mask) or None when there was no ReLU.
    dgamma / dbeta (fp32 [C]) are ACCUMULATED into when given (the parameters' .grad), else returned fresh.
    relu_beta (with y_relu None): BN + ReLU without residual - the mask is recomputed from x_raw (no read of y)."""
    require_device()
    Cc, dev = x_raw.C, x_raw.t.device
    sums = _sums(dev, 2 * Cc)
    if dx is None:
        dx = ops.new_act(x_raw.N, x_raw.H, x_raw.W, Cc, x_raw.dt, dev)
    if dgamma is None:
        dgamma = torch.zeros(Cc, dtype=torch.float32, device=dev)
        dbeta = torch.zeros(Cc, dtype=torch.float32, device=dev)
    dz = ops.new_act(x_raw.N, x_raw.H, x_raw.W, Cc, x_raw.dt, dev) if want_dz else None
    y = y_relu
    check(load().vu_bn_bwd(x_raw.t.data_ptr(), _npix(x_raw), Cc, x_raw.ctot, x_raw.coff, dy.t.data_ptr(), dy.ctot, dy.coff,
                           y.t.data_ptr() if y else None, y.ctot if y else 8, y.coff if y else 0, mean.data_ptr(),
                           invstd.data_ptr(), _f32(gamma).data_ptr(), sums.data_ptr(), dx.t.data_ptr(), dx.ctot, dx.coff,
                           dz.t.data_ptr() if dz else None, dz.ctot if dz else 8, dz.coff if dz else 0,
                           _f32(dgamma).data_ptr(), _f32(dbeta).data_ptr(), _ptr(relu_beta), x_raw.dt, _stream()), "vu_bn_bwd")
    _count(3)
    return dx, dgamma, dbeta, dz


def add_(dst: Act, src: Act) -> Act:
    require_device()
    assert (dst.N, dst.H, dst.W, dst.C) == (src.N, src.H, src.W, src.C)
    check(load().vu_add_nhwc(dst.t.data_ptr(), dst.ctot, dst.coff, src.t.data_ptr(), src.ctot, src.coff, _npix(dst), dst.C,
                             dst.dt, _stream()), "vu_add_nhwc")
    _count()
    return dst


def maxpool_bwd(x: Act, dy: Act, k: int, s: int, p: int) -> Act:
    require_device()
    dx = ops.new_act(x.N, x.H, x.W, x.C, x.dt, x.t.device)
    check(load().vu_maxpool_bwd(x.t.data_ptr(), dy.t.data_ptr(), dx.t.data_ptr(), x.N, x.H, x.W, x.C, x.ctot, x.coff, dy.ctot,
                                dy.coff, dx.ctot, dx.coff, k, s, p, x.dt, _stream()), "vu_maxpool_bwd")
    _count()
    return dx


def maxpool_idx_fwd(x: Act, k: int, s: int, p: int):
    """-> (y, code): max-pool + the uint8 position of each window's first maximum (for maxpool_idx_bwd)."""
    require_device()
    Ho, Wo = (x.H + 2 * p - k) // s + 1, (x.W + 2 * p - k) // s + 1
    y = ops.new_act(x.N, Ho, Wo, x.C, x.dt, x.t.device)
    code = torch.empty((x.N, Ho, Wo, x.C), dtype=torch.uint8, device=x.t.device)
    check(load().vu_maxpool_idx_fwd(x.t.data_ptr(), y.t.data_ptr(), code.data_ptr(), x.N, x.H, x.W, x.C, x.ctot, x.coff,
                                    y.ctot, y.coff, k, s, p, x.dt, _stream()), "vu_maxpool_idx_fwd")
    _count()
    return y, code


def maxpool_idx_bwd(code: torch.Tensor, dy: Act, H: int, W: int, k: int, s: int, p: int) -> Act:
    require_device()
    dx = ops.new_act(dy.N, H, W, dy.C, dy.dt, dy.t.device)
    check(load().vu_maxpool_idx_bwd(code.data_ptr(), dy.t.data_ptr(), dx.t.data_ptr(), dy.N, H, W, dy.C, dy.ctot, dy.coff,
                                    dx.ctot, dx.coff, k, s, p, dy.dt, _stream()), "vu_maxpool_idx_bwd")
    _count()
    return dx


def bilinear_ac_bwd(dy: Act, Hs: int, Ws: int, dx: Optional[Act] = None, accumulate: bool = False) -> Act:
    require_device()
    if dx is None:
        assert not accumulate
        dx = ops.new_act(dy.N, Hs, Ws, dy.C, dy.dt, dy.t.device)
    check(load().vu_bilinear_ac_bwd(dy.t.data_ptr(), dx.t.data_ptr(), dy.N, Hs, Ws, dy.C, dy.H, dy.W, dy.ctot, dy.coff, dx.ctot,
                                    dx.coff, int(accumulate), dy.dt, _stream()), "vu_bilinear_ac_bwd")
    _count()
    return dx


def resize_map_bwd(dy: torch.Tensor, Hs: int, Ws: int) -> torch.Tensor:
    require_device()
    dy = _f32(dy)
    N, Ho, Wo = dy.shape
    dx = torch.empty((N, Hs, Ws), dtype=torch.float32, device=dy.device)
    check(load().vu_resize_map_bwd(dy.data_ptr(), dx.data_ptr(), N, Hs, Ws, Ho, Wo, _stream()), "vu_resize_map_bwd")
    _count()
    return dx


def scale_by_map(a: Act, m: torch.Tensor, out: Act, accumulate: bool = False) -> Act:
    require_device()
    check(load().vu_scale_by_map(a.t.data_ptr(), a.ctot, a.coff, _f32(m).data_ptr(), out.t.data_ptr(), out.ctot, out.coff,
                                 _npix(a), a.C, int(accumulate), a.dt, _stream()), "vu_scale_by_map")
    _count()
    return out


def channel_dot(a: Act, b: Act) -> torch.Tensor:
    require_device()
    out = torch.empty((a.N, a.H, a.W), dtype=torch.float32, device=a.t.device)
    check(load().vu_channel_dot(a.t.data_ptr(), a.ctot, a.coff, b.t.data_ptr(), b.ctot, b.coff, out.data_ptr(), _npix(a), a.C,
                                a.dt, _stream()), "vu_channel_dot")
    _count()
    return out


def map_expand(m: torch.Tensor, w: torch.Tensor, mask: Optional[Act], out: Act) -> Act:
    require_device()
    check(load().vu_map_expand(_f32(m).data_ptr(), _f32(w).data_ptr(), mask.t.data_ptr() if mask else None,
                               mask.ctot if mask else 8, mask.coff if mask else 0, out.t.data_ptr(), out.ctot, out.coff,
                               _npix(out), out.C, out.dt, _stream()), "vu_map_expand")
    _count()
    return out


def weighted_channel_sum(x: Act, m: Optional[torch.Tensor], per_image: bool) -> torch.Tensor:
    """-> fp32 [N, C] (per_image) or [C]: sum over pixels of (m or 1) * x."""
    require_device()
    n_out = x.N if per_image else 1
    sums = torch.empty(n_out * x.C, dtype=torch.float64, device=x.t.device)
    check(load().vu_weighted_channel_sum(x.t.data_ptr(), x.ctot, x.coff, _ptr(m), x.N, x.H * x.W, x.C, int(per_image),
                                         sums.data_ptr(), x.dt, _stream()), "vu_weighted_channel_sum")
    _count(2)
    out = sums.float()
    return out.reshape(x.N, x.C) if per_image else out


def map_bn_sigmoid_fwd(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, running_mean, running_var):
    """gamma / beta: device fp32[1] (the live psi.1 parameters)."""
    require_device()
    x = _f32(x)
    sums = _sums(x.device, 2)
    stat = torch.empty(2, dtype=torch.float32, device=x.device)
    psi = torch.empty_like(x)
    check(load().vu_map_bn_sigmoid_fwd(x.data_ptr(), x.numel(), _f32(gamma).data_ptr(), _f32(beta).data_ptr(), BN_EPS, BN_MOMENTUM,
                                       _ptr(running_mean), _ptr(running_var), sums.data_ptr(), stat.data_ptr(), psi.data_ptr(),
                                       _stream()), "vu_map_bn_sigmoid_fwd")
    _count(4)
    return psi, stat


def map_bn_sigmoid_bwd(x, dpsi, psi, stat, gamma: torch.Tensor, dgamma: Optional[torch.Tensor] = None,
                       dbeta: Optional[torch.Tensor] = None):
    """-> (dx, dgamma[1], dbeta[1]); dgamma / dbeta are accumulated into when given."""
    require_device()
    sums = _sums(x.device, 2)
    dx = torch.empty_like(x)
    if dgamma is None:
        dgamma = torch.zeros(1, dtype=torch.float32, device=x.device)
        dbeta = torch.zeros(1, dtype=torch.float32, device=x.device)
    check(load().vu_map_bn_sigmoid_bwd(_f32(x).data_ptr(), _f32(dpsi).data_ptr(), _f32(psi).data_ptr(), stat.data_ptr(),
                                       _f32(gamma).data_ptr(), x.numel(), sums.data_ptr(), dx.data_ptr(),
                                       _f32(dgamma).data_ptr(), _f32(dbeta).data_ptr(), _stream()), "vu_map_bn_sigmoid_bwd")
    _count(3)
    return dx, dgamma, dbeta


def map_sum(x: torch.Tensor) -> torch.Tensor:
    require_device()
    sums = _sums(x.device, 2)
    check(load().vu_map_sum(_f32(x).data_ptr(), x.numel(), sums.data_ptr(), _stream()), "vu_map_sum")
    _count(2)
    return sums[0].float()


def vec_layer_fwd(z, w, b, gamma, beta, running_mean, running_var, relu: bool, stat_count: float):
    """z [M,L] fp32, w [C,L] -> (y, saved) with saved = (z, y_raw, y, mean, invstd)."""
    require_device()
    z = _f32(z.contiguous()); w = _f32(w)
    M, L = z.shape
    Cc = w.shape[0]
    dev = z.device
    y_raw = torch.empty((M, Cc), dtype=torch.float32, device=dev)
    y = torch.empty_like(y_raw)
    mean = torch.empty(Cc, dtype=torch.float32, device=dev) if gamma is not None else None
    invstd = torch.empty(Cc, dtype=torch.float32, device=dev) if gamma is not None else None
    check(load().vu_vec_layer_fwd(z.data_ptr(), w.data_ptr(), _ptr(b), M, L, Cc, _ptr(gamma), _ptr(beta), BN_EPS, BN_MOMENTUM,
                                  float(stat_count), _ptr(running_mean), _ptr(running_var), int(relu), y_raw.data_ptr(),
                                  y.data_ptr(), _ptr(mean), _ptr(invstd), _stream()), "vu_vec_layer_fwd")
    _count()
    return y, (z, y_raw, y, mean, invstd)


def vec_layer_bwd(dy, saved, w, gamma, relu: bool, dz: Optional[torch.Tensor], want_bias: bool, dw=None, db=None, dg=None,
                  dbeta=None):
    """-> (dw [C,L], db or None, dgamma or None, dbeta or None); dz and any gradient buffer passed in are accumulated into."""
    require_device()
    z, y_raw, y, mean, invstd = saved
    M, L = z.shape
    Cc = w.shape[0]
    dev = z.device
    zeros = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
    if dw is None:
        dw = zeros(Cc, L)
    if db is None and want_bias:
        db = zeros(Cc)
    if gamma is not None and dg is None:
        dg, dbeta = zeros(Cc), zeros(Cc)
    assert dw.numel() == Cc * L and dw.is_contiguous()
    check(load().vu_vec_layer_bwd(_f32(dy.contiguous()).data_ptr(), z.data_ptr(), _f32(w).data_ptr(), y_raw.data_ptr(),
                                  y.data_ptr(), M, L, Cc, _ptr(gamma), _ptr(mean), _ptr(invstd), int(relu), _ptr(dz),
                                  dw.data_ptr(), _ptr(db), _ptr(dg), _ptr(dbeta),
                                  torch.empty(M * Cc, dtype=torch.float32, device=dev).data_ptr() if dz is not None else None,
                                  _stream()), "vu_vec_layer_bwd")
    _count(2 if dz is not None else 1)
    return dw, db, dg, dbeta


def reparam_bwd(dz, eps, logvar, temperature, dmu, dlogvar) -> None:
    require_device()
    n = logvar.numel()
    check(load().vu_reparam_bwd(_f32(dz.contiguous()).data_ptr(), _f32(eps.contiguous()).data_ptr(), _f32(logvar).data_ptr(),
                                float(temperature), _f32(dmu).data_ptr(), _f32(dlogvar).data_ptr(), n, dz.numel() // n,
                                _stream()), "vu_reparam_bwd")
    _count()


def stem_conv_raw(x: torch.Tensor, w: torch.Tensor, dtype) -> Act:
    """conv 7x7 s2 p3 without BN / activation (train-mode BN consumes the raw output)."""
    require_device()
    x = _f32(x.float().contiguous())
    N, Cin, H, W = x.shape
    Cout = w.shape[0]
    out = ops.new_act(N, (H - 1) // 2 + 1, (W - 1) // 2 + 1, Cout, dtype, x.device)
    one = torch.ones(Cout, device=x.device); zero = torch.zeros(Cout, device=x.device)
    check(load().vu_stem_conv7x7s2_act(x.data_ptr(), _f32(w).data_ptr(), one.data_ptr(), zero.data_ptr(), out.t.data_ptr(), N,
                                       Cin, H, W, Cout, out.ctot, out.coff, 0, out.dt, _stream()), "vu_stem_conv7x7s2_act")
    _count()
    return out


def im2col_small_cin(x: torch.Tensor, k: int, stride: int, pad: int, dtype) -> Act:
    """NCHW fp32 image -> NHWC 16-bit patch matrix [N,Ho,Wo,Kp] (Kp = Cin*k*k rounded up to 8, zero padded)."""
    require_device()
    x = _f32(x.float().contiguous())
    N, Cin, H, W = x.shape
    Kp = (Cin * k * k + 7) // 8 * 8
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    out = ops.new_act(N, Ho, Wo, Kp, dtype, x.device)
    check(load().vu_im2col_small_cin(x.data_ptr(), out.t.data_ptr(), N, Cin, H, W, k, stride, pad, Kp, dtype, _stream()),
          "vu_im2col_small_cin")
    _count()
    return out


def pack_rows_padded(w: torch.Tensor, I_pad: int, dtype, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32 [O, ...] -> 16-bit [O, 1, I_pad] (a 1x1 filter bank over im2col patches)."""
    require_device()
    w = _f32(w)
    O = w.shape[0]
    I = w.numel() // O
    if out is None:
        out = torch.empty((O, 1, I_pad), dtype=ops._TORCH_DT[dtype], device=w.device)
    check(load().vu_pack_rows_padded(w.data_ptr(), out.data_ptr(), O, I, I_pad, dtype, _stream()), "vu_pack_rows_padded")
    _count()
    return out


def stem_wgrad(x: torch.Tensor, dy: Act, out) -> torch.Tensor:
    """dW [Cout, Cin, 7, 7] of the stem conv, ACCUMULATED into `out` (an fp32 tensor; an int Cout allocates zeros)."""
    require_device()
    x = _f32(x.float().contiguous())
    N, Cin, H, W = x.shape
    dw = torch.zeros((out, Cin, 7, 7), dtype=torch.float32, device=x.device) if isinstance(out, int) else _f32(out)
    Cout = dw.shape[0]
    check(load().vu_stem_wgrad(x.data_ptr(), dy.t.data_ptr(), dy.ctot, dy.coff, dw.data_ptr(), N, Cin, H, W, Cout, dy.dt,
                               _stream()), "vu_stem_wgrad")
    _count()
    return dw


def sumsq(g: torch.Tensor, out: torch.Tensor, zero_first: bool) -> None:
    require_device()
    check(load().vu_sumsq(_f32(g).data_ptr(), g.numel(), out.data_ptr(), int(zero_first), _stream()), "vu_sumsq")
    _count()


def adamw_step(p, g, m, v, lr, betas, eps, wd, step, grad_sumsq, max_norm, grad_scale=1.0) -> None:
    require_device()
    check(load().vu_adamw_step(_f32(p).data_ptr(), _f32(g).data_ptr(), _f32(m).data_ptr(), _f32(v).data_ptr(), p.numel(),
                               float(lr), float(betas[0]), float(betas[1]), float(eps), float(wd), int(step), _ptr(grad_sumsq),
                               float(max_norm), float(grad_scale), _stream()), "vu_adamw_step")
    _count()
