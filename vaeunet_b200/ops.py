"""Thin torch-tensor wrappers over the C ABI (include/vaeunet_b200.h).

PyTorch is plumbing here: it owns device memory and the current stream; every arithmetic
operation below is one of this repo's CUDA kernels. Activations are NHWC 16-bit buffers
(`Act`), possibly a channel slice of a wider concat buffer.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib
from ._lib import ConvArgs, check, load, require_device

BF16, F16 = 0, 1
_TORCH_DT = {BF16: torch.bfloat16, F16: torch.float16}
# default 16-bit storage type of activations / filter banks (see unet/_base.py KernelModule.compute_dtype)
DEFAULT_DTYPE = F16

import os as _os
S2_TAP_MASK = _os.environ.get("VU_S2_TAP_MASK", "1") != "0"    # stride-2 data gradient: skip the all-zero taps of the parity banks
# launch counter (bench.py reports it as gpu_launches)
LAUNCHES = 0
# when set to a list, conv2d appends (start_event, end_event, algorithmic_flop) per launch (bench.py roofline)
CONV_TIMING = None


def _count(n: int = 1) -> None:
    global LAUNCHES
    LAUNCHES += n


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _f32(t: torch.Tensor) -> torch.Tensor:
    assert t.dtype == torch.float32 and t.is_cuda and t.is_contiguous(), "expected contiguous CUDA fp32"
    return t


@dataclass
class Act:
    """Channel slice [coff, coff+C) of an NHWC 16-bit buffer t[N,H,W,ctot]."""
    t: torch.Tensor
    coff: int = 0
    C: int = -1

    def __post_init__(self):
        assert self.t.dim() == 4 and self.t.is_contiguous() and self.t.is_cuda
        if self.C < 0:
            self.C = self.t.shape[3] - self.coff

    @property
    def N(self): return self.t.shape[0]
    @property
    def H(self): return self.t.shape[1]
    @property
    def W(self): return self.t.shape[2]
    @property
    def ctot(self): return self.t.shape[3]
    @property
    def dt(self): return BF16 if self.t.dtype == torch.bfloat16 else F16

    def slice(self, coff: int, C: int) -> "Act":
        assert 0 <= coff and coff + C <= self.C
        return Act(self.t, self.coff + coff, C)


def new_act(N, H, W, Cc, dtype=BF16, device=None) -> Act:
    return Act(torch.empty((N, H, W, Cc), dtype=_TORCH_DT[dtype], device=device or torch.cuda.current_device()))


# ------------------------------------------------------------------ weights
def pack_conv_weight(w: torch.Tensor, dtype=BF16, transpose_io=False) -> torch.Tensor:
    """OIHW fp32 -> [O][KH*KW][I] 16-bit."""
    require_device()
    w = _f32(w.detach().float().contiguous())
    if transpose_io:
        I, O, KH, KW = w.shape
    else:
        O, I, KH, KW = w.shape
    out = torch.empty((O, KH * KW, I), dtype=_TORCH_DT[dtype], device=w.device)
    check(load().vu_pack_conv_weight(w.data_ptr(), out.data_ptr(), O, I, KH, KW, int(transpose_io), dtype, _stream()),
          "vu_pack_conv_weight")
    _count()
    return out


def pack_convT2x2_weight(w: torch.Tensor, dtype=BF16) -> torch.Tensor:
    require_device()
    w = _f32(w.detach().float().contiguous())
    I, O, KH, KW = w.shape
    assert KH == 2 and KW == 2
    out = torch.empty((4, O, 1, I), dtype=_TORCH_DT[dtype], device=w.device)
    check(load().vu_pack_convT2x2_weight(w.data_ptr(), out.data_ptr(), I, O, dtype, _stream()), "vu_pack_convT2x2_weight")
    _count()
    return out


# ------------------------------------------------------------------ conv
def conv2d(x: Act, w_packed: torch.Tensor, k: int, stride: int, pad: int, scale: Optional[torch.Tensor],
           bias: Optional[torch.Tensor], relu: bool, out: Optional[Act] = None, residual: Optional[Act] = None,
           out_up: int = 1, out_dh: int = 0, out_dw: int = 0, max_ctas: int = 0,
           fuse_w: Optional[torch.Tensor] = None, fuse_bias: float = 0.0, fuse_out: Optional[torch.Tensor] = None,
           store: bool = True, pos_bias: Optional[torch.Tensor] = None,
           stat_sums: Optional[torch.Tensor] = None, tap_mask: int = 0, x_elem_offset: int = 0) -> Optional[Act]:
    """fuse_w/fuse_bias/fuse_out: fused 1x1 Cout->1 output conv (fp32 [N,Ho,Wo]); store=False skips writing y.
    x_elem_offset: the input base pointer is advanced by this many 16-bit elements (a shifted view that keeps the image /
    row pitches of x: the (a, b) parity plane of a tensor read by a stride-2 1x1 conv starts (a*W + b)*ctot elements in).
    pos_bias: fp32 [N, 9, Cout] pre-scale bias by border class (`border_bias`): spatially constant input channels left out.
    stat_sums: ZEROED fp64 [2*Cout]; the epilogue adds per-channel sum / sum of squares of the stored outputs (the
    statistics a following train-mode BatchNorm needs: bn_train_fwd(sums=))."""
    require_device()
    Cout = w_packed.shape[0]
    assert w_packed.shape[1] == k * k and w_packed.shape[2] == x.C, (tuple(w_packed.shape), k, x.C)
    Ho = (x.H + 2 * pad - k) // stride + 1
    Wo = (x.W + 2 * pad - k) // stride + 1
    if not store:
        assert fuse_out is not None and out is None
    else:
        if out is None:
            out = new_act(x.N, Ho * out_up, Wo * out_up, Cout, x.dt, x.t.device)
        assert out.C == Cout and out.N == x.N and out.H == Ho * out_up and out.W == Wo * out_up, "conv output shape"
    a = ConvArgs()
    a.x = x.t.data_ptr() + 2 * x_elem_offset; a.N = x.N; a.H = x.H; a.W = x.W; a.x_ctot = x.ctot; a.x_coff = x.coff; a.Cin = x.C
    a.w = w_packed.data_ptr(); a.Cout = Cout; a.KH = k; a.KW = k; a.stride = stride; a.pad = pad
    a.scale = _ptr(scale); a.bias = _ptr(bias)
    if residual is not None:
        assert residual.N == x.N and residual.H == Ho and residual.W == Wo and residual.C == Cout
        a.residual = residual.t.data_ptr(); a.res_ctot = residual.ctot; a.res_coff = residual.coff
    else:
        a.residual = None
    a.relu = int(relu)
    if store:
        a.y = out.t.data_ptr(); a.y_ctot = out.ctot; a.y_coff = out.coff
    else:
        a.y = None; a.y_ctot = Cout; a.y_coff = 0
    if fuse_out is not None:
        assert fuse_out.dtype == torch.float32 and fuse_out.is_contiguous() and fuse_out.numel() == x.N * Ho * Wo
        a.fuse_w = _f32(fuse_w).data_ptr(); a.fuse_bias = float(fuse_bias); a.fuse_out = fuse_out.data_ptr()
    else:
        a.fuse_w = None; a.fuse_bias = 0.0; a.fuse_out = None
    a.out_up = out_up; a.out_dh = out_dh; a.out_dw = out_dw
    a.dtype = x.dt; a.max_ctas = max_ctas
    if pos_bias is not None:
        assert tuple(pos_bias.shape) == (x.N, 9, Cout), (tuple(pos_bias.shape), (x.N, 9, Cout))
        a.pos_bias = _f32(pos_bias).data_ptr()
    else:
        a.pos_bias = None
    if stat_sums is not None:
        assert stat_sums.dtype == torch.float64 and stat_sums.numel() == 2 * Cout and stat_sums.is_contiguous() and store
        a.stat_sums = stat_sums.data_ptr()
    else:
        a.stat_sums = None
    a.tap_mask = tap_mask
    if CONV_TIMING is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(load().vu_conv2d_fwd(C.byref(a), _stream()), "vu_conv2d_fwd")
        e1.record()
        ntaps = bin(tap_mask).count("1") if tap_mask else k * k
        CONV_TIMING.append((e0, e1, 2.0 * x.N * Ho * Wo * Cout * ntaps * x.C,
                            f"N{x.N} {x.H}x{x.W} Cin{x.C} Cout{Cout} k{k} s{stride}" + (f" taps{ntaps}" if tap_mask else "")))
    else:
        check(load().vu_conv2d_fwd(C.byref(a), _stream()), "vu_conv2d_fwd")
    _count()
    return out


def conv2d_wgrad(x: Act, dy: Act, k: int, stride: int, pad: int, out: Optional[torch.Tensor] = None,
                 oihw: bool = False, x_elem_offset: int = 0) -> torch.Tensor:
    """dW of conv2d(x, W), fp32, ACCUMULATED into `out` if given: packed tap order [Cout, k*k, Cin], or (oihw) the
    parameter's own layout [Cout, Cin, k, k]."""
    require_device()
    Cout = dy.C
    Ho = (x.H + 2 * pad - k) // stride + 1
    Wo = (x.W + 2 * pad - k) // stride + 1
    assert (dy.N, dy.H, dy.W) == (x.N, Ho, Wo) and dy.dt == x.dt, "wgrad: dy shape"
    shape = (Cout, x.C, k, k) if oihw else (Cout, k * k, x.C)
    if out is None:
        out = torch.zeros(shape, dtype=torch.float32, device=x.t.device)
    assert tuple(out.shape) == shape and out.dtype == torch.float32 and out.is_contiguous()
    a = ConvArgs()
    a.x = x.t.data_ptr() + 2 * x_elem_offset; a.N = x.N; a.H = x.H; a.W = x.W; a.x_ctot = x.ctot; a.x_coff = x.coff; a.Cin = x.C
    a.Cout = Cout; a.KH = k; a.KW = k; a.stride = stride; a.pad = pad; a.dtype = x.dt
    fn = load().vu_conv2d_wgrad_oihw if oihw else load().vu_conv2d_wgrad
    if CONV_TIMING is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(fn(C.byref(a), dy.t.data_ptr(), dy.ctot, dy.coff, out.data_ptr(), _stream()), "vu_conv2d_wgrad")
        e1.record()
        CONV_TIMING.append((e0, e1, 2.0 * x.N * Ho * Wo * Cout * k * k * x.C,
                            f"wgrad N{x.N} {x.H}x{x.W} Cin{x.C} Cout{Cout} k{k} s{stride}"))
    else:
        check(fn(C.byref(a), dy.t.data_ptr(), dy.ctot, dy.coff, out.data_ptr(), _stream()), "vu_conv2d_wgrad")
    _count()
    return out


def unpack_wgrad(dw_packed: torch.Tensor, k: int) -> torch.Tensor:
    """[Cout, k*k, Cin] -> OIHW."""
    O, T, I = dw_packed.shape
    return dw_packed.reshape(O, k, k, I).permute(0, 3, 1, 2).contiguous()


def pack_dgrad_weights(w: torch.Tensor, stride: int, dtype=BF16, out: Optional[torch.Tensor] = None):
    """Filter banks for the data gradient of conv2d(x, w) (w OIHW fp32, 3x3 p1 or 1x1 p0), run by vu_conv2d_fwd on dY
    (vu_pack_dgrad_weight): stride 1 -> one bank (w rotated by 180 degrees, in/out swapped); stride 2 -> one bank per
    output-pixel parity (ph, pc) of dX, written with the pixel-shuffle store (out_up = 2). Returns the list of banks
    (views of one [nbanks, I, k*k, O] tensor, `out` if given)."""
    require_device()
    w = _f32(w.detach())
    O, I, k, _ = w.shape
    nb = 4 if (stride == 2 and k == 3) else 1
    if out is None:
        out = torch.empty((nb, I, k * k, O), dtype=_TORCH_DT[dtype], device=w.device)
    assert tuple(out.shape) == (nb, I, k * k, O)
    check(load().vu_pack_dgrad_weight(w.data_ptr(), out.data_ptr(), O, I, k, stride, dtype, _stream()), "vu_pack_dgrad_weight")
    _count()
    return [out[b] for b in range(nb)]


def conv2d_dgrad(dy: Act, banks, k: int, stride: int, H: int, W: int, out: Optional[Act] = None,
                 accumulate: bool = False) -> Act:
    """dX [N,H,W,Cin] of conv2d(x, w) from dY; `banks` from pack_dgrad_weights. accumulate: dX += (stride 1 only)."""
    require_device()
    Cin = banks[0].shape[0]
    if out is None:
        assert not accumulate
        out = new_act(dy.N, H, W, Cin, dy.dt, dy.t.device)
    if stride == 1:
        assert (dy.H, dy.W) == (H, W)
        return conv2d(dy, banks[0], k, 1, k // 2, None, None, False, out=out, residual=out if accumulate else None)
    assert not accumulate and H == 2 * dy.H and W == 2 * dy.W, "stride-2 dgrad needs even input sizes"
    if k == 1:
        out.t[..., out.coff:out.coff + out.C].zero_()
        _count()
        return conv2d(dy, banks[0], 1, 1, 0, None, None, False, out=out, out_up=2, out_dh=0, out_dw=0)
    # non-zero bank rows / columns per parity (vu_pack_dgrad_weight): parity 0 -> {1}, parity 1 -> {1, 2}; the other taps of a
    # parity bank are identically zero and are skipped (9 tap passes over the four banks instead of 36)
    nz = ((1,), (1, 2))
    for ph in range(2):
        for pc in range(2):
            mask = 0
            for r in nz[ph]:
                for s in nz[pc]:
                    mask |= 1 << (r * 3 + s)
            conv2d(dy, banks[ph * 2 + pc], 3, 1, 1, None, None, False, out=out, out_up=2, out_dh=ph, out_dw=pc,
                   tap_mask=mask if S2_TAP_MASK else 0)
    return out


def border_bias(w_oihw: torch.Tensor, coff: int, zvec: torch.Tensor) -> torch.Tensor:
    """-> fp32 [D, 9, Cout]: exact 3x3 p1 conv contribution of the spatially constant channels [coff, coff+Cz) with per-image
    values zvec [D, Cz], per border class of the output pixel (vu_border_bias); feeds conv2d(pos_bias=)."""
    require_device()
    w = _f32(w_oihw)
    Cout, Cin_total, kh, kw = w.shape
    assert kh == 3 and kw == 3
    zvec = _f32(zvec.contiguous())
    D, Cz = zvec.shape
    out = torch.empty((D, 9, Cout), dtype=torch.float32, device=w.device)
    check(load().vu_border_bias(w.data_ptr(), Cout, Cin_total, coff, Cz, zvec.data_ptr(), D, out.data_ptr(), _stream()),
          "vu_border_bias")
    _count()
    return out


def border_wsum(w_oihw: torch.Tensor, coff: int, Cz: int) -> torch.Tensor:
    """-> fp32 [9, Cout, Cz]: per border class, the sum of the in-image taps of w[:, coff:coff+Cz] (built once per weight)."""
    require_device()
    w = _f32(w_oihw)
    Cout, Cin_total = w.shape[:2]
    out = torch.empty((9, Cout, Cz), dtype=torch.float32, device=w.device)
    check(load().vu_border_wsum(w.data_ptr(), Cout, Cin_total, coff, Cz, out.data_ptr(), _stream()), "vu_border_wsum")
    _count()
    return out


def border_bias_apply(wsum: torch.Tensor, zvec: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """wsum [9, Cout, Cz], zvec [D, Cz] -> pos_bias fp32 [D, 9, Cout] (added to `out` when given)."""
    require_device()
    zvec = _f32(zvec.contiguous())
    _, Cout, Cz = wsum.shape
    D = zvec.shape[0]
    assert zvec.shape[1] == Cz
    acc = out is not None
    if out is None:
        out = torch.empty((D, 9, Cout), dtype=torch.float32, device=zvec.device)
    assert tuple(out.shape) == (D, 9, Cout)
    check(load().vu_border_bias_apply(_f32(wsum).data_ptr(), zvec.data_ptr(), D, Cout, Cz, _f32(out).data_ptr(), int(acc), _stream()),
          "vu_border_bias_apply")
    _count()
    return out


# ------------------------------------------------------------------ layout / pooling / resize
def nchw_to_nhwc(x: torch.Tensor, out: Optional[Act] = None, dtype=BF16) -> Act:
    require_device()
    x = _f32(x.float().contiguous())
    N, Cc, H, W = x.shape
    if out is None:
        out = new_act(N, H, W, Cc, dtype, x.device)
    check(load().vu_nchw_to_nhwc(x.data_ptr(), out.t.data_ptr(), N, Cc, H, W, out.ctot, out.coff, out.dt, _stream()),
          "vu_nchw_to_nhwc")
    _count()
    return out


def nhwc_to_nchw(x: Act) -> torch.Tensor:
    require_device()
    y = torch.empty((x.N, x.C, x.H, x.W), dtype=torch.float32, device=x.t.device)
    check(load().vu_nhwc_to_nchw(x.t.data_ptr(), y.data_ptr(), x.N, x.C, x.H, x.W, x.ctot, x.coff, x.dt, _stream()),
          "vu_nhwc_to_nchw")
    _count()
    return y


def stem_conv7x7s2(x: torch.Tensor, w: torch.Tensor, scale: torch.Tensor, bias: torch.Tensor, dtype=BF16,
                   out: Optional[Act] = None) -> Act:
    require_device()
    x = _f32(x.float().contiguous())
    N, Cin, H, W = x.shape
    Cout = w.shape[0]
    Ho, Wo = (H + 6 - 7) // 2 + 1, (W + 6 - 7) // 2 + 1
    if out is None:
        out = new_act(N, Ho, Wo, Cout, dtype, x.device)
    check(load().vu_stem_conv7x7s2(x.data_ptr(), _f32(w).data_ptr(), _f32(scale).data_ptr(), _f32(bias).data_ptr(),
                                   out.t.data_ptr(), N, Cin, H, W, Cout, out.ctot, out.coff, out.dt, _stream()),
          "vu_stem_conv7x7s2")
    _count()
    return out


def conv3x3_small_cin(x: torch.Tensor, w: torch.Tensor, scale: torch.Tensor, bias: torch.Tensor, relu: bool,
                      dtype=BF16, out: Optional[Act] = None) -> Act:
    require_device()
    x = _f32(x.float().contiguous())
    N, Cin, H, W = x.shape
    Cout = w.shape[0]
    if out is None:
        out = new_act(N, H, W, Cout, dtype, x.device)
    check(load().vu_conv3x3_small_cin(x.data_ptr(), _f32(w).data_ptr(), _f32(scale).data_ptr(), _f32(bias).data_ptr(),
                                      out.t.data_ptr(), N, Cin, H, W, Cout, out.ctot, out.coff, int(relu), out.dt,
                                      _stream()), "vu_conv3x3_small_cin")
    _count()
    return out


def maxpool(x: Act, k: int, s: int, p: int, out: Optional[Act] = None) -> Act:
    require_device()
    Ho, Wo = (x.H + 2 * p - k) // s + 1, (x.W + 2 * p - k) // s + 1
    if out is None:
        out = new_act(x.N, Ho, Wo, x.C, x.dt, x.t.device)
    check(load().vu_maxpool_nhwc(x.t.data_ptr(), out.t.data_ptr(), x.N, x.H, x.W, x.C, x.ctot, x.coff, k, s, p,
                                 out.ctot, out.coff, x.dt, _stream()), "vu_maxpool_nhwc")
    _count()
    return out


def bilinear_ac(x: Act, Ho: int, Wo: int, out: Optional[Act] = None, pad_t: int = 0, pad_l: int = 0) -> Act:
    require_device()
    if out is None:
        out = new_act(x.N, Ho, Wo, x.C, x.dt, x.t.device)
    check(load().vu_bilinear_ac_nhwc(x.t.data_ptr(), out.t.data_ptr(), x.N, x.H, x.W, x.C, x.ctot, x.coff, Ho, Wo,
                                     out.H, out.W, pad_t, pad_l, out.ctot, out.coff, x.dt, _stream()),
          "vu_bilinear_ac_nhwc")
    _count()
    return out


def broadcast_vec(v: torch.Tensor, out: Act) -> Act:
    """v fp32 [N, C] -> constant map in out[N,H,W,slice]."""
    require_device()
    v = _f32(v)
    assert v.shape == (out.N, out.C)
    check(load().vu_broadcast_vec_nhwc(v.data_ptr(), out.t.data_ptr(), out.N, out.H, out.W, out.C, out.ctot, out.coff,
                                       out.dt, _stream()), "vu_broadcast_vec_nhwc")
    _count()
    return out


# ------------------------------------------------------------------ attention gate / latent / head
def attention_gate(g1: Act, x1: Act, skip: Act, psi_w: torch.Tensor, psi_scale: float, psi_bias: float, out: Act,
                   psi_out: Optional[torch.Tensor] = None) -> Act:
    require_device()
    Ng, Ns = g1.N, skip.N
    assert Ng % Ns == 0 and x1.N == Ns and g1.C == x1.C and out.C == skip.C and out.N == Ng
    assert (g1.H, g1.W) == (skip.H, skip.W) == (x1.H, x1.W) == (out.H, out.W)
    check(load().vu_attention_gate(g1.t.data_ptr(), g1.ctot, g1.coff, x1.t.data_ptr(), x1.ctot, x1.coff,
                                   skip.t.data_ptr(), skip.ctot, skip.coff, _f32(psi_w).data_ptr(), float(psi_scale),
                                   float(psi_bias), out.t.data_ptr(), out.ctot, out.coff, _ptr(psi_out), Ng, Ng // Ns,
                                   g1.H, g1.W, g1.C, skip.C, g1.dt, _stream()), "vu_attention_gate")
    _count()
    return out


def decoder_concat(x_low: Act, g1_low: Act, x1: Act, skip: Act, psi_w: torch.Tensor, psi_scale: float, psi_bias: float,
                   zvec: Optional[torch.Tensor], out: Act, psi_out: Optional[torch.Tensor] = None) -> Act:
    """out[D,H,W,:] = [bilinear_ac(x_low) | skip * psi | zvec] with psi from bilinear_ac(g1_low) + x1 (one pass)."""
    require_device()
    D, H, W = out.N, out.H, out.W
    Dp = skip.N
    assert D % Dp == 0 and x1.N == Dp and x_low.N == D and g1_low.N == D and out.coff == 0
    assert (x_low.H, x_low.W) == (g1_low.H, g1_low.W) and (skip.H, skip.W) == (x1.H, x1.W) == (H, W)
    assert g1_low.C == x1.C == psi_w.numel()
    cz = 0
    if zvec is not None:
        zvec = _f32(zvec)
        assert zvec.shape[0] == D
        cz = zvec.shape[1]
    assert x_low.C + skip.C + cz == out.C, (x_low.C, skip.C, cz, out.C)
    check(load().vu_decoder_concat(x_low.t.data_ptr(), x_low.H, x_low.W, x_low.C, x_low.ctot, x_low.coff,
                                   g1_low.t.data_ptr(), g1_low.ctot, g1_low.coff, x1.t.data_ptr(), x1.ctot, x1.coff,
                                   skip.t.data_ptr(), skip.ctot, skip.coff, g1_low.C, skip.C, _f32(psi_w).data_ptr(),
                                   float(psi_scale), float(psi_bias), _ptr(zvec), cz, out.t.data_ptr(), out.ctot,
                                   _ptr(psi_out), D, D // Dp, H, W, out.dt, _stream()), "vu_decoder_concat")
    _count()
    return out


def latent_head(f4: Act, w_mu, b_mu, w_lv, b_lv, eps: Optional[torch.Tensor], temperature: float, Ns: int):
    """-> mu [B,L], logvar [B,L], z [Ns,B,L] (fp32)."""
    require_device()
    B, L = f4.N, w_mu.shape[0]
    dev = f4.t.device
    mu = torch.empty((B, L), dtype=torch.float32, device=dev)
    lv = torch.empty((B, L), dtype=torch.float32, device=dev)
    z = torch.empty((max(Ns, 1), B, L), dtype=torch.float32, device=dev) if Ns > 0 else None
    if eps is not None:
        eps = _f32(eps)
        assert eps.shape == (Ns, B, L), (tuple(eps.shape), (Ns, B, L))
    check(load().vu_latent_head(f4.t.data_ptr(), B, f4.H, f4.W, f4.C, f4.ctot, f4.coff, _f32(w_mu).data_ptr(),
                                _f32(b_mu).data_ptr(), _f32(w_lv).data_ptr(), _f32(b_lv).data_ptr(), _ptr(eps),
                                float(temperature), Ns, L, mu.data_ptr(), lv.data_ptr(), _ptr(z), f4.dt, _stream()),
          "vu_latent_head")
    _count()
    return mu, lv, z


def reparameterize(mu: torch.Tensor, logvar: torch.Tensor, eps: torch.Tensor, temperature: float = 1.0) -> torch.Tensor:
    """z = mu + eps * exp(.5 logvar) * T ; eps [..., *mu.shape] (leading sample dims broadcast over mu)."""
    require_device()
    mu = _f32(mu.contiguous()); logvar = _f32(logvar.contiguous()); eps = _f32(eps.contiguous())
    n = mu.numel()
    assert eps.numel() % n == 0 and eps.shape[-mu.dim():] == mu.shape
    z = torch.empty_like(eps)
    check(load().vu_reparameterize(mu.data_ptr(), logvar.data_ptr(), eps.data_ptr(), float(temperature), z.data_ptr(),
                                   n, eps.numel() // n, _stream()), "vu_reparameterize")
    _count()
    return z


def latent_project(z: torch.Tensor, w: torch.Tensor, scale: Optional[torch.Tensor], shift: Optional[torch.Tensor],
                   relu: bool) -> torch.Tensor:
    require_device()
    z = _f32(z)
    M, L = z.shape
    Cc = w.shape[0]
    v = torch.empty((M, Cc), dtype=torch.float32, device=z.device)
    check(load().vu_latent_project(z.data_ptr(), _f32(w).data_ptr(), _ptr(scale), _ptr(shift), v.data_ptr(), M, L, Cc,
                                   int(relu), _stream()), "vu_latent_project")
    _count()
    return v


def final_conv1(x: Act, w: torch.Tensor, bias: float, out: Optional[torch.Tensor] = None,
                bias_dev: Optional[torch.Tensor] = None) -> torch.Tensor:
    """bias_dev: device fp32[1] added to `bias` (training reads the live parameter without a host sync)."""
    require_device()
    if out is None:
        out = torch.empty((x.N, x.H, x.W), dtype=torch.float32, device=x.t.device)
    check(load().vu_final_conv1(x.t.data_ptr(), x.N, x.H, x.W, x.C, x.ctot, x.coff, _f32(w).data_ptr(), float(bias),
                                _ptr(bias_dev), out.data_ptr(), x.dt, _stream()), "vu_final_conv1")
    _count()
    return out


SIG_NONE, SIG_PRE, SIG_POST = 0, 1, 2


def resize_logits(x: torch.Tensor, Ho: int, Wo: int, sigmoid_mode: int = SIG_NONE) -> torch.Tensor:
    """x fp32 [N,Hs,Ws] -> [N,Ho,Wo], bilinear align_corners=True."""
    require_device()
    x = _f32(x)
    N, Hs, Ws = x.shape
    y = torch.empty((N, Ho, Wo), dtype=torch.float32, device=x.device)
    check(load().vu_resize_logits(x.data_ptr(), y.data_ptr(), N, Hs, Ws, Ho, Wo, sigmoid_mode, _stream()),
          "vu_resize_logits")
    _count()
    return y


def uncertainty_reduce(x: torch.Tensor, Ho: int, Wo: int, input_is_prob: bool = False, want_probs: bool = False,
                       outs=("mean", "std", "entropy", "mutual_info", "coeff_var")):
    """x fp32 [B,Ns,Hs,Ws] (logits, or probabilities at output size) -> dict of [B,Ho,Wo] maps."""
    require_device()
    x = _f32(x)
    B, Ns, Hs, Ws = x.shape
    dev = x.device
    res = {k: torch.empty((B, Ho, Wo), dtype=torch.float32, device=dev) for k in outs}
    probs = torch.empty((B, Ns, Ho, Wo), dtype=torch.float32, device=dev) if want_probs else None
    g = lambda k: _ptr(res.get(k))
    check(load().vu_uncertainty_reduce(x.data_ptr(), B, Ns, Hs, Ws, Ho, Wo, int(input_is_prob), g("mean"), g("std"),
                                       g("entropy"), g("mutual_info"), g("coeff_var"), _ptr(probs), _stream()),
          "vu_uncertainty_reduce")
    _count()
    if want_probs:
        res["samples"] = probs
    return res


def moments_partial(probs: torch.Tensor):
    """probs fp32 [Ns, ...] -> (mean, M2, sum_entropy) per pixel."""
    require_device()
    probs = _f32(probs)
    Ns = probs.shape[0]
    shp = probs.shape[1:]
    npix = int(probs[0].numel()) if Ns > 0 else 0
    mk = lambda: torch.zeros(shp, dtype=torch.float32, device=probs.device)
    mean, m2, sh = mk(), mk(), mk()
    if Ns > 0:
        check(load().vu_moments_partial(probs.data_ptr(), Ns, npix, mean.data_ptr(), m2.data_ptr(), sh.data_ptr(),
                                        _stream()), "vu_moments_partial")
        _count()
    return mean, m2, sh


def moments_finalize(counts, mean: torch.Tensor, m2: torch.Tensor, sh: torch.Tensor):
    """Chan-merge stacked shard moments [S, ...] with per-shard sample counts -> dict of maps."""
    require_device()
    S = mean.shape[0]
    shp = mean.shape[1:]
    npix = int(mean[0].numel())
    arr = (C.c_float * S)(*[float(c) for c in counts])
    res = {k: torch.empty(shp, dtype=torch.float32, device=mean.device)
           for k in ("mean", "std", "entropy", "mutual_info", "coeff_var")}
    check(load().vu_moments_finalize(C.cast(arr, C.c_void_p), _f32(mean).data_ptr(), _f32(m2).data_ptr(),
                                     _f32(sh).data_ptr(), S, npix, res["mean"].data_ptr(), res["std"].data_ptr(),
                                     res["entropy"].data_ptr(), res["mutual_info"].data_ptr(),
                                     res["coeff_var"].data_ptr(), _stream()), "vu_moments_finalize")
    _count()
    return res


def blend_accumulate(tile: torch.Tensor, sh: int, sw: int, ti: int, tj: int, nth: int, ntw: int, overlap: int,
                     acc: torch.Tensor, wsum: torch.Tensor) -> None:
    require_device()
    tile = _f32(tile)
    ph, pw = tile.shape[-2:]
    H, W = acc.shape[-2:]
    check(load().vu_blend_accumulate(tile.data_ptr(), ph, pw, sh, sw, ti, tj, nth, ntw, overlap, _f32(acc).data_ptr(),
                                     _f32(wsum).data_ptr(), H, W, _stream()), "vu_blend_accumulate")
    _count()


def blend_finalize(acc: torch.Tensor, wsum: torch.Tensor) -> torch.Tensor:
    require_device()
    out = torch.empty_like(acc)
    check(load().vu_blend_finalize(_f32(acc).data_ptr(), _f32(wsum).data_ptr(), out.data_ptr(), acc.numel(), _stream()),
          "vu_blend_finalize")
    _count()
    return out


# ------------------------------------------------------------------ losses / metrics
def seg_loss_sums(x: torch.Tensor, target: torch.Tensor, input_is_prob=False, focal_alpha=0.8, focal_gamma=2.0):
    """-> fp64[8]: sum p, sum t, sum pt, sum bce, sum focal, count, sum (p-t)^2, sum nll."""
    require_device()
    x = _f32(x.contiguous()); target = _f32(target.float().contiguous())
    assert x.numel() == target.numel()
    sums = torch.empty(8, dtype=torch.float64, device=x.device)
    check(load().vu_seg_loss_sums(x.data_ptr(), target.data_ptr(), x.numel(), int(input_is_prob), float(focal_alpha),
                                  float(focal_gamma), sums.data_ptr(), _stream()), "vu_seg_loss_sums")
    _count()
    return sums


def seg_loss_bwd(x, target, sums, w_bce, w_dice, w_focal, focal_alpha=0.8, focal_gamma=2.0, smooth=1.0,
                 grad_scale=1.0) -> torch.Tensor:
    require_device()
    x = _f32(x.contiguous()); target = _f32(target.float().contiguous())
    dx = torch.empty_like(x)
    check(load().vu_seg_loss_bwd(x.data_ptr(), target.data_ptr(), x.numel(), sums.data_ptr(), float(w_bce),
                                 float(w_dice), float(w_focal), float(focal_alpha), float(focal_gamma), float(smooth),
                                 float(grad_scale), dx.data_ptr(), _stream()), "vu_seg_loss_bwd")
    _count()
    return dx


def kl_free_bits(mu: torch.Tensor, logvar: torch.Tensor, free_bits: float, want_grad=False, grad_scale=1.0):
    require_device()
    mu = _f32(mu.contiguous()); logvar = _f32(logvar.contiguous())
    B, L = mu.shape
    kl = torch.empty((), dtype=torch.float32, device=mu.device)
    dmu = torch.empty_like(mu) if want_grad else None
    dlv = torch.empty_like(mu) if want_grad else None
    check(load().vu_kl_free_bits(mu.data_ptr(), logvar.data_ptr(), B, L, float(free_bits), kl.data_ptr(), _ptr(dmu),
                                 _ptr(dlv), float(grad_scale), _stream()), "vu_kl_free_bits")
    _count()
    return (kl, dmu, dlv) if want_grad else kl


def ece_hist(p: torch.Tensor, gt: torch.Tensor, num_bins: int = 10) -> torch.Tensor:
    """-> fp64 [3, bins]: count, sum gt, sum p."""
    require_device()
    p = _f32(p.contiguous()); gt = _f32(gt.float().contiguous())
    hist = torch.empty((3, num_bins), dtype=torch.float64, device=p.device)
    check(load().vu_ece_hist(p.data_ptr(), gt.data_ptr(), p.numel(), num_bins, hist.data_ptr(), _stream()), "vu_ece_hist")
    _count(2)
    return hist


def confusion_counts(x: torch.Tensor, t: torch.Tensor, thr: float = 0.5) -> torch.Tensor:
    """-> fp64 [4]: tp, fp, tn, fn of (x > thr) against (t > 0.5)."""
    require_device()
    x = _f32(x.contiguous()); t = _f32(t.float().contiguous())
    c = torch.empty(4, dtype=torch.float64, device=x.device)
    check(load().vu_confusion_counts(x.data_ptr(), t.data_ptr(), x.numel(), float(thr), c.data_ptr(), _stream()),
          "vu_confusion_counts")
    _count(2)
    return c


# ------------------------------------------------------------------ sort-based uncertainty metrics
def _sort_ws(n: int, dev) -> torch.Tensor:
    return torch.empty(int(load().vu_sort_workspace_bytes(n)), dtype=torch.uint8, device=dev)


def sort_pairs_desc(score: torch.Tensor, vals: torch.Tensor, ws: Optional[torch.Tensor] = None):
    """Stable sort of (score fp32, payload 32-bit) pairs by score DESCENDING. -> (keys uint32-as-int32: order-preserving
    keys whose equality == score equality, payload in sorted order, workspace)."""
    require_device()
    score = _f32(score.contiguous().view(-1))
    n = score.numel()
    assert vals.numel() == n and vals.element_size() == 4 and vals.is_contiguous()
    if ws is None:
        ws = _sort_ws(n, score.device)
    keys = torch.empty(n, dtype=torch.int32, device=score.device)
    out = torch.empty(n, dtype=vals.dtype, device=score.device)
    check(load().vu_sort_pairs_desc_f32(score.data_ptr(), vals.data_ptr(), n, keys.data_ptr(), out.data_ptr(), ws.data_ptr(),
                                        ws.numel(), _stream()), "vu_sort_pairs_desc_f32")
    _count(1 + 3 * 8)
    return keys, out, ws


def pixel_nll_errors(p: torch.Tensor, gt: torch.Tensor, eps: float = 1e-7):
    """-> (err fp32 like p, stats dict: sum, nan_count, min, max) — uncertainty_metrics.py:179-181."""
    require_device()
    p = _f32(p.contiguous()); gt = _f32(gt.float().contiguous())
    err = torch.empty_like(p)
    stats = torch.zeros(4, dtype=torch.float64, device=p.device)
    check(load().vu_pixel_nll_errors(p.data_ptr(), gt.data_ptr(), p.numel(), float(eps), err.data_ptr(), stats.data_ptr(),
                                     _stream()), "vu_pixel_nll_errors")
    _count()
    return err, stats


def decode_nll_stats(stats: torch.Tensor):
    """host view of pixel_nll_errors' statistics (one device -> host read)."""
    import struct
    raw = stats.cpu()
    s, nans = float(raw[0]), float(raw[1])
    mm = raw[2:3].view(torch.int32)
    dec = lambda u: struct.unpack("<f", struct.pack("<I", (u ^ 0x80000000) if (u >> 31) else (~u & 0xFFFFFFFF)))[0]
    lo, hi = int(mm[0]) & 0xFFFFFFFF, int(mm[1]) & 0xFFFFFFFF
    return s, nans, dec(lo), dec(hi)


def sorted_bin_sums(vals_sorted: torch.Tensor, cuts: torch.Tensor) -> torch.Tensor:
    """vals_sorted: fp32 values (any 32-bit dtype holding float bits); cuts int64 [nb+1] on the device -> fp64 [nb]."""
    require_device()
    nb = cuts.numel() - 1
    assert cuts.dtype == torch.int64 and cuts.is_cuda and cuts.is_contiguous()
    sums = torch.empty(nb, dtype=torch.float64, device=vals_sorted.device)
    check(load().vu_sorted_bin_sums(vals_sorted.data_ptr(), vals_sorted.numel(), cuts.data_ptr(), nb, sums.data_ptr(), _stream()),
          "vu_sorted_bin_sums")
    _count()
    return sums


def masked_sum(x: torch.Tensor, bits: torch.Tensor) -> torch.Tensor:
    """sum of fp32 x over the set bits of a packed uint8 keep-mask (little bit order) -> fp64 scalar tensor."""
    require_device()
    x = _f32(x.contiguous())
    assert bits.dtype == torch.uint8 and bits.is_cuda and bits.numel() * 8 >= x.numel()
    out = torch.empty(1, dtype=torch.float64, device=x.device)
    check(load().vu_masked_sum(x.data_ptr(), bits.data_ptr(), x.numel(), out.data_ptr(), _stream()), "vu_masked_sum")
    _count()
    return out


def error_labels(p: torch.Tensor, gt: torch.Tensor, thr: float = 0.5) -> torch.Tensor:
    require_device()
    p = _f32(p.contiguous()); gt = _f32(gt.float().contiguous())
    lab = torch.empty(p.numel(), dtype=torch.int32, device=p.device)
    check(load().vu_error_labels(p.data_ptr(), gt.data_ptr(), p.numel(), float(thr), lab.data_ptr(), _stream()), "vu_error_labels")
    _count()
    return lab


def binary_curve_auc(keys_sorted: torch.Tensor, labels_sorted: torch.Tensor, ws: torch.Tensor) -> torch.Tensor:
    """-> fp64 [4]: un-normalised ROC area, un-normalised average precision, positives, negatives."""
    require_device()
    out = torch.empty(4, dtype=torch.float64, device=keys_sorted.device)
    check(load().vu_binary_curve_auc(keys_sorted.data_ptr(), labels_sorted.data_ptr(), keys_sorted.numel(), ws.data_ptr(),
                                     ws.numel(), out.data_ptr(), _stream()), "vu_binary_curve_auc")
    _count(5)
    return out


# ------------------------------------------------------------------ patch pipeline / batched blending
def tile_gather(src: torch.Tensor, coords: torch.Tensor, ph: int, pw: int, divisor: float = 255.0,
                binarise: bool = False) -> torch.Tensor:
    """Tiles of ONE device-resident image -> fp32 NCHW batch [T, C, ph, pw]. src: uint8 HWC [H, W, C] (divided by `divisor`)
    or fp32 CHW [C, H, W] / [1, C, H, W]; coords int32 [T, 2] = (y0, x0) on the device."""
    require_device()
    assert coords.dtype == torch.int32 and coords.is_cuda and coords.is_contiguous() and coords.shape[1] == 2
    T = coords.shape[0]
    if src.dtype == torch.uint8:
        assert src.dim() == 3 and src.is_contiguous()
        H, W, Cc = src.shape
        is_u8 = 1
    else:
        src = _f32(src.contiguous())
        if src.dim() == 4:
            assert src.shape[0] == 1
            src = src[0]
        Cc, H, W = src.shape
        is_u8 = 0
    out = torch.empty((T, Cc, ph, pw), dtype=torch.float32, device=src.device)
    check(load().vu_tile_gather(src.data_ptr(), is_u8, Cc, H, W, coords.data_ptr(), T, ph, pw, float(divisor), int(binarise),
                                out.data_ptr(), _stream()), "vu_tile_gather")
    _count()
    return out


def tile_stats(tiles: torch.Tensor, masks: Optional[torch.Tensor], black_thr: float = 0.1) -> torch.Tensor:
    """-> fp32 [T, 2]: black-border ratio, lesion flag (data_loading.py:287-300,382)."""
    require_device()
    tiles = _f32(tiles)
    T, Cc, ph, pw = tiles.shape
    if masks is not None:
        masks = _f32(masks.contiguous())
        assert masks.numel() == T * ph * pw
    out = torch.empty((T, 2), dtype=torch.float32, device=tiles.device)
    check(load().vu_tile_stats(tiles.data_ptr(), T, Cc, ph, pw, _ptr(masks), float(black_thr), out.data_ptr(), _stream()),
          "vu_tile_stats")
    _count()
    return out


def resample_u8(src: torch.Tensor, out_size: int, axis: int, bounds: torch.Tensor, kk: torch.Tensor) -> torch.Tensor:
    """One pass of Pillow's 8-bit resampling (axis 0 = x, 1 = y) of a uint8 HWC image with host-built tables."""
    require_device()
    assert src.dtype == torch.uint8 and src.dim() == 3 and src.is_contiguous()
    Hs, Ws, Cc = src.shape
    Hd, Wd = (Hs, out_size) if axis == 0 else (out_size, Ws)
    assert bounds.dtype == torch.int32 and kk.dtype == torch.int32 and bounds.shape == (out_size, 2) and kk.shape[0] == out_size
    dst = torch.empty((Hd, Wd, Cc), dtype=torch.uint8, device=src.device)
    check(load().vu_resample_u8(src.data_ptr(), Hs, Ws, Cc, dst.data_ptr(), Hd, Wd, axis, bounds.contiguous().data_ptr(),
                                kk.contiguous().data_ptr(), kk.shape[1], _stream()), "vu_resample_u8")
    _count()
    return dst


def nearest_u8(src: torch.Tensor, yi: torch.Tensor, xi: torch.Tensor) -> torch.Tensor:
    require_device()
    assert src.dtype == torch.uint8 and src.dim() == 3 and src.is_contiguous()
    Hs, Ws, Cc = src.shape
    dst = torch.empty((yi.numel(), xi.numel(), Cc), dtype=torch.uint8, device=src.device)
    check(load().vu_nearest_u8(src.data_ptr(), Hs, Ws, Cc, dst.data_ptr(), yi.numel(), xi.numel(), yi.data_ptr(), xi.data_ptr(),
                               _stream()), "vu_nearest_u8")
    _count()
    return dst


def blend_accumulate_batch(tiles: torch.Tensor, info: torch.Tensor, bbox, nth: int, ntw: int, overlap: int, acc: torch.Tensor,
                           wsum: torch.Tensor) -> None:
    """tiles fp32 [T, Ns, ph, pw]; info int32 [T, 4] = (sh, sw, ti, tj) on the device; bbox = (y0, y1, x0, x1) host ints;
    acc [Ns, H, W]; wsum [H, W] (one weight map for all samples)."""
    require_device()
    tiles = _f32(tiles)
    T, Ns, ph, pw = tiles.shape
    H, W = acc.shape[-2:]
    assert acc.shape[0] == Ns and tuple(wsum.shape) == (H, W) and info.dtype == torch.int32 and info.shape == (T, 4)
    y0, y1, x0, x1 = bbox
    check(load().vu_blend_accumulate_batch(tiles.data_ptr(), info.data_ptr(), T, Ns, ph, pw, nth, ntw, overlap, _f32(acc).data_ptr(),
                                           _f32(wsum).data_ptr(), H, W, y0, y1, x0, x1, _stream()), "vu_blend_accumulate_batch")
    _count()


def blend_finalize_shared(acc: torch.Tensor, wsum: torch.Tensor) -> torch.Tensor:
    """acc [Ns, H, W] / (wsum [H, W] + 1e-8)."""
    require_device()
    out = torch.empty_like(acc)
    plane = wsum.numel()
    check(load().vu_blend_finalize_shared(_f32(acc).data_ptr(), _f32(wsum).data_ptr(), out.data_ptr(), plane, acc.numel() // plane,
                                          _stream()), "vu_blend_finalize_shared")
    _count()
    return out
