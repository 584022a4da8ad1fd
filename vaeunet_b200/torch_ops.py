"""torch custom ops over the C ABI: `torch.ops.vaeunet.*` (BASELINE north_star "these become torch custom ops over a
thin C-ABI extension").

Every op is a `torch.library.custom_op` whose implementation enqueues this repo's CUDA kernels through the ctypes
binding (ops.py / train_ops.py -> libvaeunet_b200.so), with
  * `register_fake`   - shape / dtype propagation without a device (torch.compile / FakeTensor tracing), and
  * `register_autograd` - the backward, itself expressed in `torch.ops.vaeunet.*` calls,
so a model built from them runs under the autograd engine: `model.train(); seg, mu, logvar = model(images);
loss.backward(); optimizer.step()` - the reference's literal training loop (train.py:381-414) - works on the drop-in
modules for EVERY configuration (UNetResNet with any latent_injection / use_skip / use_attention, the basic UNet with
ConvTranspose2d or bilinear up-sampling). `vaeunet_b200.training.TrainStep` remains the fused fast path for the default
configuration.

Tensor conventions: activations are plain contiguous torch tensors [N, H, W, C] in the 16-bit compute dtype (NHWC);
parameters are the modules' own fp32 nn.Parameters (OIHW conv weights); maps are fp32 [N, H, W]. There is no CPU
implementation: the ops raise VaeUnetError without a B200.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
from torch import Tensor

from . import ops
from . import train_ops as T
from .ops import Act

_lib_op = torch.library.custom_op
NS = "vaeunet"


def _dt(t: Tensor) -> int:
    return ops.BF16 if t.dtype == torch.bfloat16 else ops.F16


def _tdt(code: int):
    return ops._TORCH_DT[code]


def _c(t: Tensor) -> Tensor:
    return t if t.is_contiguous() else t.contiguous()


def _out_hw(H: int, W: int, k: int, stride: int, pad: int) -> Tuple[int, int]:
    return (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1


# ---------------------------------------------------------------------------------------------------- convolution
@_lib_op(f"{NS}::conv2d", mutates_args=())
def conv2d(x: Tensor, w: Tensor, bias: Optional[Tensor], stride: int, pad: int) -> Tensor:
    """nn.Conv2d (3x3 p1 / 1x1 p0, stride 1|2) on the tcgen05 implicit-GEMM kernels. x NHWC 16-bit, w OIHW fp32."""
    k = w.shape[2]
    wp = ops.pack_conv_weight(w, _dt(x))
    return ops.conv2d(Act(_c(x)), wp, k, stride, pad, None, bias, False).t


@conv2d.register_fake
def _(x, w, bias, stride, pad):
    N, H, W_, _ = x.shape
    Ho, Wo = _out_hw(H, W_, w.shape[2], stride, pad)
    return x.new_empty((N, Ho, Wo, w.shape[0]))


@_lib_op(f"{NS}::conv2d_dgrad", mutates_args=())
def conv2d_dgrad(dy: Tensor, w: Tensor, stride: int, H: int, W: int) -> Tensor:
    """dX of conv2d(x, w): the forward kernels run on dY with rotated / parity filter banks (vu_pack_dgrad_weight)."""
    k = w.shape[2]
    banks = ops.pack_dgrad_weights(w, stride, _dt(dy))
    return ops.conv2d_dgrad(Act(_c(dy)), banks, k, stride, H, W).t


@conv2d_dgrad.register_fake
def _(dy, w, stride, H, W):
    return dy.new_empty((dy.shape[0], H, W, w.shape[1]))


@_lib_op(f"{NS}::conv2d_wgrad", mutates_args=())
def conv2d_wgrad(x: Tensor, dy: Tensor, k: int, stride: int, pad: int) -> Tensor:
    """dW (OIHW fp32) of conv2d(x, w): MN-major tcgen05 GEMM over the pixels."""
    dw = ops.conv2d_wgrad(Act(_c(x)), Act(_c(dy)), k, stride, pad)            # packed [O][k*k][I]
    return ops.unpack_wgrad(dw, k)


@conv2d_wgrad.register_fake
def _(x, dy, k, stride, pad):
    return x.new_empty((dy.shape[3], x.shape[3], k, k), dtype=torch.float32)


@_lib_op(f"{NS}::channel_sum", mutates_args=())
def channel_sum(x: Tensor) -> Tensor:
    """sum over N, H, W of an NHWC 16-bit tensor -> fp32 [C] (bias gradients)."""
    return T.weighted_channel_sum(Act(_c(x)), None, per_image=False)


@channel_sum.register_fake
def _(x):
    return x.new_empty((x.shape[3],), dtype=torch.float32)


def _conv2d_setup(ctx, inputs, output):
    x, w, bias, stride, pad = inputs
    ctx.save_for_backward(x, w)
    ctx.stride, ctx.pad, ctx.has_bias = stride, pad, bias is not None


def _conv2d_bwd(ctx, dy):
    x, w = ctx.saved_tensors
    dy = _c(dy)
    dx = dw = db = None
    if ctx.needs_input_grad[0]:
        dx = torch.ops.vaeunet.conv2d_dgrad(dy, w, ctx.stride, x.shape[1], x.shape[2])
    if ctx.needs_input_grad[1]:
        dw = torch.ops.vaeunet.conv2d_wgrad(x, dy, w.shape[2], ctx.stride, ctx.pad)
    if ctx.has_bias and ctx.needs_input_grad[2]:
        db = torch.ops.vaeunet.channel_sum(dy)
    return dx, dw, db, None, None


conv2d.register_autograd(_conv2d_bwd, setup_context=_conv2d_setup)


# ---- first conv on the NCHW fp32 image (Cin <= 4): im2col patches + 1x1 tensor-core GEMM
@_lib_op(f"{NS}::conv_small_cin", mutates_args=())
def conv_small_cin(images: Tensor, w: Tensor, stride: int, pad: int, dtype: int) -> Tensor:
    """ResNet stem 7x7 s2 p3 / basic-UNet `inc` 3x3 p1 on the raw image: NCHW fp32 -> NHWC 16-bit (no bias, BN follows)."""
    k = w.shape[2]
    xcol = T.im2col_small_cin(images, k, stride, pad, dtype)
    wp = T.pack_rows_padded(w, xcol.C, dtype)
    return ops.conv2d(xcol, wp, 1, 1, 0, None, None, False).t


@conv_small_cin.register_fake
def _(images, w, stride, pad, dtype):
    N, _, H, W_ = images.shape
    Ho, Wo = _out_hw(H, W_, w.shape[2], stride, pad)
    return images.new_empty((N, Ho, Wo, w.shape[0]), dtype=_tdt(dtype))


@_lib_op(f"{NS}::conv_small_cin_wgrad", mutates_args=())
def conv_small_cin_wgrad(images: Tensor, dy: Tensor, k: int, stride: int, pad: int) -> Tensor:
    xcol = T.im2col_small_cin(images, k, stride, pad, _dt(dy))                # recomputed: the patch matrix is not kept
    dw = ops.conv2d_wgrad(xcol, Act(_c(dy)), 1, 1, 0)                         # [O][1][Kp]
    cin = images.shape[1]
    return dw.view(dw.shape[0], -1)[:, :cin * k * k].reshape(dw.shape[0], cin, k, k).contiguous()


@conv_small_cin_wgrad.register_fake
def _(images, dy, k, stride, pad):
    return images.new_empty((dy.shape[3], images.shape[1], k, k), dtype=torch.float32)


def _csc_setup(ctx, inputs, output):
    images, w, stride, pad, dtype = inputs
    ctx.save_for_backward(images)
    ctx.k, ctx.stride, ctx.pad = w.shape[2], stride, pad


def _csc_bwd(ctx, dy):
    (images,) = ctx.saved_tensors
    dw = torch.ops.vaeunet.conv_small_cin_wgrad(images, _c(dy), ctx.k, ctx.stride, ctx.pad)
    return None, dw, None, None, None            # no gradient w.r.t. the input image


conv_small_cin.register_autograd(_csc_bwd, setup_context=_csc_setup)


# ---- ConvTranspose2d k2 s2 (unet/unet_parts.py:76): four 1x1 GEMMs with a pixel-shuffle store
@_lib_op(f"{NS}::conv_transpose2x2", mutates_args=())
def conv_transpose2x2(x: Tensor, w: Tensor, bias: Optional[Tensor]) -> Tensor:
    """x [N,H,W,I], w [I,O,2,2] fp32 -> [N,2H,2W,O]."""
    dt = _dt(x)
    wp = ops.pack_convT2x2_weight(w, dt)
    xa = Act(_c(x))
    out = ops.new_act(xa.N, 2 * xa.H, 2 * xa.W, w.shape[1], dt, x.device)
    for ab in range(4):
        ops.conv2d(xa, wp[ab], 1, 1, 0, None, bias, False, out=out, out_up=2, out_dh=ab // 2, out_dw=ab % 2)
    return out.t


@conv_transpose2x2.register_fake
def _(x, w, bias):
    return x.new_empty((x.shape[0], 2 * x.shape[1], 2 * x.shape[2], w.shape[1]))


@_lib_op(f"{NS}::conv_transpose2x2_bwd", mutates_args=())
def conv_transpose2x2_bwd(x: Tensor, w: Tensor, dy: Tensor) -> Tuple[Tensor, Tensor]:
    """-> (dX [N,H,W,I], dW [I,O,2,2]). dY[n, 2h+a, 2w+b, :] is the (a, b) parity plane of dY, i.e. a stride-2 1x1 view of
    the tensor whose base is shifted by (a, b) pixels: dX = sum_ab conv1x1_s2(dY_ab, W_ab) (accumulated through the
    residual input of the epilogue) and dW_ab = wgrad1x1_s2(dY_ab as the input, x as the output gradient)."""
    dt = _dt(x)
    I, O = w.shape[0], w.shape[1]
    xa, dya = Act(_c(x)), Act(_c(dy))
    dx = ops.new_act(xa.N, xa.H, xa.W, I, dt, x.device)
    dw = torch.empty((2, 2, I, O), dtype=torch.float32, device=x.device)
    wt = w.permute(2, 3, 0, 1).contiguous()                      # [a][b][I][O] fp32 (layout shuffle of a tiny tensor)
    for a in range(2):
        for b in range(2):
            off = (a * dya.W + b) * dya.ctot                     # element offset of the (a, b) parity plane's origin
            wp = ops.pack_conv_weight(wt[a, b].reshape(I, O, 1, 1), dt)           # [I][1][O]
            first = (a == 0 and b == 0)
            ops.conv2d(dya, wp, 1, 2, 0, None, None, False, out=dx, residual=None if first else dx, x_elem_offset=off)
            g = ops.conv2d_wgrad(dya, xa, 1, 2, 0, x_elem_offset=off)             # [Cout'=I][1][Cin'=O]
            dw[a, b] = g.view(I, O)
    return dx.t, dw.permute(2, 3, 0, 1).contiguous()


@conv_transpose2x2_bwd.register_fake
def _(x, w, dy):
    return x.new_empty(x.shape), w.new_empty(w.shape)


def _ct_setup(ctx, inputs, output):
    x, w, bias = inputs
    ctx.save_for_backward(x, w)
    ctx.has_bias = bias is not None


def _ct_bwd(ctx, dy):
    x, w = ctx.saved_tensors
    dy = _c(dy)
    dx, dw = torch.ops.vaeunet.conv_transpose2x2_bwd(x, w, dy)
    db = torch.ops.vaeunet.channel_sum(dy) if ctx.has_bias else None
    return dx, dw, db


conv_transpose2x2.register_autograd(_ct_bwd, setup_context=_ct_setup)


# ---------------------------------------------------------------------------------------------------- BatchNorm (train mode)
@_lib_op(f"{NS}::batch_norm_train", mutates_args=())
def batch_norm_train(x: Tensor, gamma: Tensor, beta: Tensor, residual: Optional[Tensor], relu: bool) -> Tuple[Tensor, Tensor, Tensor]:
    """y = act(BN_batch-stats(x) + residual) on an NHWC 16-bit map -> (y, mean, invstd). Functional: the module's running
    statistics are updated by vaeunet::bn_update_running from (mean, invstd)."""
    y, mean, invstd = T.bn_train_fwd(Act(_c(x)), gamma, beta, None, None, relu,
                                     Act(_c(residual)) if residual is not None else None, update_running=False)
    return y.t, mean, invstd


@batch_norm_train.register_fake
def _(x, gamma, beta, residual, relu):
    C = x.shape[3]
    return x.new_empty(x.shape), x.new_empty((C,), dtype=torch.float32), x.new_empty((C,), dtype=torch.float32)


@_lib_op(f"{NS}::bn_update_running", mutates_args=("running_mean", "running_var"))
def bn_update_running(running_mean: Tensor, running_var: Tensor, mean: Tensor, invstd: Tensor, count: float) -> None:
    """running_mean / running_var of nn.BatchNorm2d (momentum 0.1, unbiased variance over `count` values per channel)."""
    T.bn_update_running(mean, invstd, count, running_mean, running_var)


@_lib_op(f"{NS}::batch_norm_train_bwd", mutates_args=())
def batch_norm_train_bwd(x: Tensor, dy: Tensor, y: Optional[Tensor], mean: Tensor, invstd: Tensor, gamma: Tensor,
                         want_dz: bool) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """-> (dx, dgamma, dbeta, dz). y: the post-ReLU output (mask) or None; dz = masked dy (the residual branch's gradient;
    an empty tensor when want_dz is False)."""
    dx, dg, db, dz = T.bn_train_bwd(Act(_c(x)), Act(_c(dy)), Act(_c(y)) if y is not None else None, mean, invstd, gamma,
                                    want_dz=want_dz)
    return dx.t, dg, db, (dz.t if dz is not None else x.new_empty((0,)))


@batch_norm_train_bwd.register_fake
def _(x, dy, y, mean, invstd, gamma, want_dz):
    C = x.shape[3]
    f = lambda: x.new_empty((C,), dtype=torch.float32)
    return x.new_empty(x.shape), f(), f(), (x.new_empty(x.shape) if want_dz else x.new_empty((0,)))


def _bn_setup(ctx, inputs, output):
    x, gamma, beta, residual, relu = inputs
    y, mean, invstd = output
    ctx.relu, ctx.has_res = relu, residual is not None
    ctx.save_for_backward(x, y if relu else None, mean, invstd, gamma)
    ctx.set_materialize_grads(False)


def _bn_bwd(ctx, dy, dmean, dinvstd):
    x, y, mean, invstd, gamma = ctx.saved_tensors
    if dy is None:
        return (None,) * 5
    dx, dg, db, dz = torch.ops.vaeunet.batch_norm_train_bwd(x, _c(dy), y, mean, invstd, gamma, ctx.has_res)
    return dx, dg, db, (dz if ctx.has_res else None), None


batch_norm_train.register_autograd(_bn_bwd, setup_context=_bn_setup)


# ---------------------------------------------------------------------------------------------------- pooling / resize
@_lib_op(f"{NS}::maxpool", mutates_args=())
def maxpool(x: Tensor, k: int, stride: int, pad: int) -> Tuple[Tensor, Tensor]:
    """MaxPool2d on NHWC 16-bit -> (y, code): code = uint8 window position of each first maximum (for the backward)."""
    y, code = T.maxpool_idx_fwd(Act(_c(x)), k, stride, pad)
    return y.t, code


@maxpool.register_fake
def _(x, k, stride, pad):
    N, H, W_, C = x.shape
    Ho, Wo = _out_hw(H, W_, k, stride, pad)
    return x.new_empty((N, Ho, Wo, C)), x.new_empty((N, Ho, Wo, C), dtype=torch.uint8)


@_lib_op(f"{NS}::maxpool_bwd", mutates_args=())
def maxpool_bwd(code: Tensor, dy: Tensor, H: int, W: int, k: int, stride: int, pad: int) -> Tensor:
    return T.maxpool_idx_bwd(code, Act(_c(dy)), H, W, k, stride, pad).t


@maxpool_bwd.register_fake
def _(code, dy, H, W, k, stride, pad):
    return dy.new_empty((dy.shape[0], H, W, dy.shape[3]))


def _mp_setup(ctx, inputs, output):
    x, k, stride, pad = inputs
    ctx.save_for_backward(output[1])
    ctx.geom = (x.shape[1], x.shape[2], k, stride, pad)
    ctx.set_materialize_grads(False)


def _mp_bwd(ctx, dy, dcode):
    (code,) = ctx.saved_tensors
    H, W_, k, stride, pad = ctx.geom
    return torch.ops.vaeunet.maxpool_bwd(code, _c(dy), H, W_, k, stride, pad), None, None, None


maxpool.register_autograd(_mp_bwd, setup_context=_mp_setup)


@_lib_op(f"{NS}::bilinear_ac", mutates_args=())
def bilinear_ac(x: Tensor, Ho: int, Wo: int) -> Tensor:
    """F.interpolate(mode='bilinear', align_corners=True) on NHWC 16-bit."""
    return ops.bilinear_ac(Act(_c(x)), Ho, Wo).t


@bilinear_ac.register_fake
def _(x, Ho, Wo):
    return x.new_empty((x.shape[0], Ho, Wo, x.shape[3]))


@_lib_op(f"{NS}::bilinear_ac_bwd", mutates_args=())
def bilinear_ac_bwd(dy: Tensor, Hs: int, Ws: int) -> Tensor:
    return T.bilinear_ac_bwd(Act(_c(dy)), Hs, Ws).t


@bilinear_ac_bwd.register_fake
def _(dy, Hs, Ws):
    return dy.new_empty((dy.shape[0], Hs, Ws, dy.shape[3]))


def _bl_setup(ctx, inputs, output):
    ctx.hw = (inputs[0].shape[1], inputs[0].shape[2])


def _bl_bwd(ctx, dy):
    return torch.ops.vaeunet.bilinear_ac_bwd(_c(dy), ctx.hw[0], ctx.hw[1]), None, None


bilinear_ac.register_autograd(_bl_bwd, setup_context=_bl_setup)


@_lib_op(f"{NS}::resize_map", mutates_args=())
def resize_map(m: Tensor, Ho: int, Wo: int) -> Tensor:
    """bilinear (align_corners=True) resize of fp32 maps [N,h,w] -> [N,Ho,Wo] (the final up-sampling of the logits)."""
    return ops.resize_logits(_c(m), Ho, Wo, ops.SIG_NONE)


@resize_map.register_fake
def _(m, Ho, Wo):
    return m.new_empty((m.shape[0], Ho, Wo))


@_lib_op(f"{NS}::resize_map_bwd", mutates_args=())
def resize_map_bwd(dy: Tensor, Hs: int, Ws: int) -> Tensor:
    return T.resize_map_bwd(_c(dy), Hs, Ws)


@resize_map_bwd.register_fake
def _(dy, Hs, Ws):
    return dy.new_empty((dy.shape[0], Hs, Ws))


def _rm_setup(ctx, inputs, output):
    ctx.hw = (inputs[0].shape[1], inputs[0].shape[2])


def _rm_bwd(ctx, dy):
    return torch.ops.vaeunet.resize_map_bwd(_c(dy), ctx.hw[0], ctx.hw[1]), None, None


resize_map.register_autograd(_rm_bwd, setup_context=_rm_setup)


# ---------------------------------------------------------------------------------------------------- attention-gate pieces
@_lib_op(f"{NS}::conv1x1_to_map", mutates_args=())
def conv1x1_to_map(x: Tensor, w: Tensor, bias: Tensor) -> Tensor:
    """1x1 conv C -> 1 (+ bias) on NHWC 16-bit -> fp32 map [N,H,W]: the gate's psi conv and final_conv / OutConv."""
    return ops.final_conv1(Act(_c(x)), w.reshape(-1), 0.0, bias_dev=bias.reshape(-1))


@conv1x1_to_map.register_fake
def _(x, w, bias):
    return x.new_empty(x.shape[:3], dtype=torch.float32)


@_lib_op(f"{NS}::conv1x1_to_map_bwd", mutates_args=())
def conv1x1_to_map_bwd(x: Tensor, w: Tensor, dm: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    xa = Act(_c(x))
    dm = _c(dm)
    dx = T.map_expand(dm, w.reshape(-1), None, ops.new_act(xa.N, xa.H, xa.W, xa.C, xa.dt, x.device))
    dw = T.weighted_channel_sum(xa, dm, per_image=False)
    db = T.map_sum(dm).reshape(1)
    return dx.t, dw, db


@conv1x1_to_map_bwd.register_fake
def _(x, w, dm):
    return x.new_empty(x.shape), x.new_empty((x.shape[3],), dtype=torch.float32), x.new_empty((1,), dtype=torch.float32)


def _c1_setup(ctx, inputs, output):
    x, w, bias = inputs
    ctx.save_for_backward(x, w)
    ctx.wshape, ctx.bshape = w.shape, bias.shape


def _c1_bwd(ctx, dm):
    x, w = ctx.saved_tensors
    dx, dw, db = torch.ops.vaeunet.conv1x1_to_map_bwd(x, w, _c(dm))
    return dx, dw.reshape(ctx.wshape), db.reshape(ctx.bshape)


conv1x1_to_map.register_autograd(_c1_bwd, setup_context=_c1_setup)


@_lib_op(f"{NS}::map_bn_sigmoid", mutates_args=())
def map_bn_sigmoid(m: Tensor, gamma: Tensor, beta: Tensor) -> Tuple[Tensor, Tensor]:
    """psi = sigmoid(BatchNorm2d(1)(m)) in train mode on an fp32 map -> (psi, stat = {mean, invstd}). Functional."""
    return T.map_bn_sigmoid_fwd(_c(m), gamma, beta, None, None)


@map_bn_sigmoid.register_fake
def _(m, gamma, beta):
    return m.new_empty(m.shape), m.new_empty((2,))


@_lib_op(f"{NS}::map_bn_sigmoid_bwd", mutates_args=())
def map_bn_sigmoid_bwd(m: Tensor, dpsi: Tensor, psi: Tensor, stat: Tensor, gamma: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    return T.map_bn_sigmoid_bwd(m, _c(dpsi), psi, stat, gamma)


@map_bn_sigmoid_bwd.register_fake
def _(m, dpsi, psi, stat, gamma):
    return m.new_empty(m.shape), m.new_empty((1,)), m.new_empty((1,))


def _ms_setup(ctx, inputs, output):
    m, gamma = inputs[0], inputs[1]
    ctx.save_for_backward(m, output[0], output[1], gamma)
    ctx.set_materialize_grads(False)


def _ms_bwd(ctx, dpsi, dstat):
    m, psi, stat, gamma = ctx.saved_tensors
    dm, dg, db = torch.ops.vaeunet.map_bn_sigmoid_bwd(m, _c(dpsi), psi, stat, gamma)
    return dm, dg.reshape(gamma.shape), db.reshape(gamma.shape)


map_bn_sigmoid.register_autograd(_ms_bwd, setup_context=_ms_setup)


@_lib_op(f"{NS}::scale_by_map", mutates_args=())
def scale_by_map(x: Tensor, m: Tensor) -> Tensor:
    """y[n,h,w,c] = x[n,h,w,c] * m[n,h,w] (skip * psi)."""
    xa = Act(_c(x))
    return T.scale_by_map(xa, _c(m), ops.new_act(xa.N, xa.H, xa.W, xa.C, xa.dt, x.device)).t


@scale_by_map.register_fake
def _(x, m):
    return x.new_empty(x.shape)


@_lib_op(f"{NS}::channel_dot", mutates_args=())
def channel_dot(a: Tensor, b: Tensor) -> Tensor:
    return T.channel_dot(Act(_c(a)), Act(_c(b)))


@channel_dot.register_fake
def _(a, b):
    return a.new_empty(a.shape[:3], dtype=torch.float32)


def _sm_setup(ctx, inputs, output):
    ctx.save_for_backward(inputs[0], inputs[1])


def _sm_bwd(ctx, dy):
    x, m = ctx.saved_tensors
    dy = _c(dy)
    return torch.ops.vaeunet.scale_by_map(dy, m), torch.ops.vaeunet.channel_dot(dy, x)


scale_by_map.register_autograd(_sm_bwd, setup_context=_sm_setup)


# ---------------------------------------------------------------------------------------------------- latent path
@_lib_op(f"{NS}::latent_head", mutates_args=())
def latent_head(f4: Tensor, w_mu: Tensor, b_mu: Tensor, w_lv: Tensor, b_lv: Tensor) -> Tuple[Tensor, Tensor]:
    """mu / logvar = GAP(conv1x1(f4) + b) (unet_resnet.py:140-147,205-206). f4 NHWC 16-bit; weights [L, C, 1, 1]."""
    L = w_mu.shape[0]
    mu, lv, _ = ops.latent_head(Act(_c(f4)), w_mu.reshape(L, -1), b_mu, w_lv.reshape(L, -1), b_lv, None, 1.0, 0)
    return mu, lv


@latent_head.register_fake
def _(f4, w_mu, b_mu, w_lv, b_lv):
    B, L = f4.shape[0], w_mu.shape[0]
    return f4.new_empty((B, L), dtype=torch.float32), f4.new_empty((B, L), dtype=torch.float32)


@_lib_op(f"{NS}::latent_head_bwd", mutates_args=())
def latent_head_bwd(f4: Tensor, w_mu: Tensor, w_lv: Tensor, dmu: Tensor, dlv: Tensor) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """-> (df4, dw_mu, db_mu, dw_lv, db_lv)."""
    fa = Act(_c(f4))
    B, L = dmu.shape
    hw = float(fa.H * fa.W)
    gap = T.weighted_channel_sum(fa, None, per_image=True) / hw
    dgap = torch.zeros_like(gap)
    grads = []
    for w, dh in ((w_mu, _c(dmu)), (w_lv, _c(dlv))):
        dw, db, _, _ = T.vec_layer_bwd(dh, (gap, dh, dh, None, None), w.reshape(L, -1), None, False, dgap, True)
        grads += [dw.reshape(w.shape), db]
    df4 = ops.broadcast_vec(dgap / hw, ops.new_act(B, fa.H, fa.W, fa.C, fa.dt, f4.device))
    return df4.t, grads[0], grads[1], grads[2], grads[3]


@latent_head_bwd.register_fake
def _(f4, w_mu, w_lv, dmu, dlv):
    L = w_mu.shape[0]
    f = lambda s: f4.new_empty(s, dtype=torch.float32)
    return f4.new_empty(f4.shape), f(w_mu.shape), f((L,)), f(w_lv.shape), f((L,))


def _lh_setup(ctx, inputs, output):
    f4, w_mu, b_mu, w_lv, b_lv = inputs
    ctx.save_for_backward(f4, w_mu, w_lv)
    ctx.set_materialize_grads(True)


def _lh_bwd(ctx, dmu, dlv):
    f4, w_mu, w_lv = ctx.saved_tensors
    df4, dwm, dbm, dwl, dbl = torch.ops.vaeunet.latent_head_bwd(f4, w_mu, w_lv, _c(dmu), _c(dlv))
    return df4, dwm, dbm, dwl, dbl


latent_head.register_autograd(_lh_bwd, setup_context=_lh_setup)


@_lib_op(f"{NS}::reparameterize", mutates_args=())
def reparameterize(mu: Tensor, logvar: Tensor, eps: Tensor, temperature: float) -> Tensor:
    """z = mu + eps * exp(0.5 logvar) * T (unet_resnet.py:191-194; utils/vae_utils.py:5-10)."""
    return ops.reparameterize(mu, logvar, eps, temperature)


@reparameterize.register_fake
def _(mu, logvar, eps, temperature):
    return eps.new_empty(eps.shape)


@_lib_op(f"{NS}::reparameterize_bwd", mutates_args=())
def reparameterize_bwd(dz: Tensor, eps: Tensor, logvar: Tensor, temperature: float) -> Tuple[Tensor, Tensor]:
    dmu, dlv = torch.zeros_like(logvar), torch.zeros_like(logvar)
    T.reparam_bwd(_c(dz), eps, logvar, temperature, dmu, dlv)
    return dmu, dlv


@reparameterize_bwd.register_fake
def _(dz, eps, logvar, temperature):
    return logvar.new_empty(logvar.shape), logvar.new_empty(logvar.shape)


def _rp_setup(ctx, inputs, output):
    mu, logvar, eps, temperature = inputs
    ctx.save_for_backward(eps, logvar)
    ctx.T = temperature


def _rp_bwd(ctx, dz):
    eps, logvar = ctx.saved_tensors
    dmu, dlv = torch.ops.vaeunet.reparameterize_bwd(_c(dz), eps, logvar, ctx.T)
    return dmu, dlv, None, None


reparameterize.register_autograd(_rp_bwd, setup_context=_rp_setup)


@_lib_op(f"{NS}::vec_layer", mutates_args=())
def vec_layer(z: Tensor, w: Tensor, b: Tensor, gamma: Tensor, beta: Tensor, relu: bool) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """1x1 conv + train-mode BatchNorm (+ReLU) on a SPATIALLY CONSTANT map = a dense layer on the per-image vectors
    z [M, L] (z_initial, z_proj, W_g of a constant map) -> (y, y_raw, mean, invstd): the statistics over the M
    rows equal those over the M*H*W identical pixels. Functional (running statistics: vaeunet::bn_update_running)."""
    y, (_, y_raw, _, mean, invstd) = T.vec_layer_fwd(_c(z), w.reshape(w.shape[0], -1), b, gamma, beta, None, None, relu, 1.0)
    return y, y_raw, mean, invstd


@vec_layer.register_fake
def _(z, w, b, gamma, beta, relu):
    M, C = z.shape[0], w.shape[0]
    return z.new_empty((M, C)), z.new_empty((M, C)), z.new_empty((C,)), z.new_empty((C,))


@_lib_op(f"{NS}::vec_layer_bwd", mutates_args=())
def vec_layer_bwd(dy: Tensor, z: Tensor, w: Tensor, y_raw: Tensor, y: Tensor, mean: Tensor, invstd: Tensor, gamma: Tensor,
                  relu: bool) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """-> (dz, dw, db, dgamma, dbeta)."""
    dz = torch.zeros_like(z)
    dw, db, dg, dbeta = T.vec_layer_bwd(_c(dy), (z, y_raw, y, mean, invstd), w.reshape(w.shape[0], -1), gamma, relu, dz, True)
    return dz, dw.reshape(w.shape), db, dg, dbeta


@vec_layer_bwd.register_fake
def _(dy, z, w, y_raw, y, mean, invstd, gamma, relu):
    C = w.shape[0]
    return z.new_empty(z.shape), w.new_empty(w.shape), z.new_empty((C,)), z.new_empty((C,)), z.new_empty((C,))


def _vl_setup(ctx, inputs, output):
    z, w, b, gamma = inputs[0], inputs[1], inputs[2], inputs[3]
    y, y_raw, mean, invstd = output
    ctx.save_for_backward(z, w, y_raw, y, mean, invstd, gamma)
    ctx.relu = inputs[5]
    ctx.set_materialize_grads(False)


def _vl_bwd(ctx, dy, dy_raw, dmean, dinvstd):
    z, w, y_raw, y, mean, invstd, gamma = ctx.saved_tensors
    dz, dw, db, dg, dbeta = torch.ops.vaeunet.vec_layer_bwd(_c(dy), z, w, y_raw, y, mean, invstd, gamma, ctx.relu)
    return dz, dw, db, dg, dbeta, None


vec_layer.register_autograd(_vl_bwd, setup_context=_vl_setup)


@_lib_op(f"{NS}::broadcast_vec", mutates_args=())
def broadcast_vec(v: Tensor, H: int, W: int, dtype: int) -> Tensor:
    """per-image vectors [N, C] fp32 -> constant NHWC 16-bit map [N, H, W, C]."""
    v = _c(v)
    return ops.broadcast_vec(v, ops.new_act(v.shape[0], H, W, v.shape[1], dtype, v.device)).t


@broadcast_vec.register_fake
def _(v, H, W, dtype):
    return v.new_empty((v.shape[0], H, W, v.shape[1]), dtype=_tdt(dtype))


@_lib_op(f"{NS}::image_channel_sum", mutates_args=())
def image_channel_sum(x: Tensor) -> Tensor:
    """sum over H, W per image -> fp32 [N, C] (adjoint of broadcast_vec)."""
    return T.weighted_channel_sum(Act(_c(x)), None, per_image=True)


@image_channel_sum.register_fake
def _(x):
    return x.new_empty((x.shape[0], x.shape[3]), dtype=torch.float32)


def _bv_bwd(ctx, dy):
    return torch.ops.vaeunet.image_channel_sum(_c(dy)), None, None, None


broadcast_vec.register_autograd(_bv_bwd, setup_context=lambda ctx, inputs, output: None)


# ---------------------------------------------------------------------------------------------------- inference entry points
@_lib_op(f"{NS}::uncertainty_reduce", mutates_args=())
def uncertainty_reduce(logits: Tensor, Ho: int, Wo: int) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """calculate_uncertainty_metrics (visualize_vae.py:90-117) fused with the final up-sampling + sigmoid: logits fp32
    [B, Ns, h, w] -> (mean, std, entropy, mutual_info, coeff_var), each [B, Ho, Wo]."""
    r = ops.uncertainty_reduce(_c(logits), Ho, Wo)
    return r["mean"], r["std"], r["entropy"], r["mutual_info"], r["coeff_var"]


@uncertainty_reduce.register_fake
def _(logits, Ho, Wo):
    f = lambda: logits.new_empty((logits.shape[0], Ho, Wo))
    return f(), f(), f(), f(), f()


OP_NAMES: List[str] = [
    "conv2d", "conv2d_dgrad", "conv2d_wgrad", "channel_sum", "conv_small_cin", "conv_small_cin_wgrad", "conv_transpose2x2",
    "conv_transpose2x2_bwd", "batch_norm_train", "batch_norm_train_bwd", "bn_update_running", "maxpool", "maxpool_bwd", "bilinear_ac",
    "bilinear_ac_bwd", "resize_map", "resize_map_bwd", "conv1x1_to_map", "conv1x1_to_map_bwd", "map_bn_sigmoid",
    "map_bn_sigmoid_bwd", "scale_by_map", "channel_dot", "latent_head", "latent_head_bwd", "reparameterize",
    "reparameterize_bwd", "vec_layer", "vec_layer_bwd", "broadcast_vec", "image_channel_sum", "uncertainty_reduce",
]
