"""ctypes binding of libvaeunet_b200.so (the C ABI declared in include/vaeunet_b200.h).

There is NO CPU fallback: importing works anywhere (so `state_dict` surgery and host logic can be
tested on CPU), but any compute call raises `VaeUnetError` unless the extension is built and a
B200 (sm_100) device is current.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "libvaeunet_b200.so"


class VaeUnetError(RuntimeError):
    pass


class ConvArgs(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("N", C.c_int), ("H", C.c_int), ("W", C.c_int),
        ("x_ctot", C.c_int), ("x_coff", C.c_int), ("Cin", C.c_int),
        ("w", C.c_void_p),
        ("Cout", C.c_int), ("KH", C.c_int), ("KW", C.c_int), ("stride", C.c_int), ("pad", C.c_int),
        ("scale", C.c_void_p), ("bias", C.c_void_p),
        ("residual", C.c_void_p), ("res_ctot", C.c_int), ("res_coff", C.c_int),
        ("relu", C.c_int),
        ("y", C.c_void_p), ("y_ctot", C.c_int), ("y_coff", C.c_int),
        ("out_up", C.c_int), ("out_dh", C.c_int), ("out_dw", C.c_int),
        ("dtype", C.c_int), ("max_ctas", C.c_int),
        ("fuse_w", C.c_void_p), ("fuse_bias", C.c_float), ("fuse_out", C.c_void_p),
        ("pos_bias", C.c_void_p),
        ("stat_sums", C.c_void_p),
        ("tap_mask", C.c_int),
    ]


i, f, p, sz = C.c_int, C.c_float, C.c_void_p, C.c_size_t
# name -> argtypes (every symbol declared in include/vaeunet_b200.h)
SIGNATURES = {
    "vu_version": [],
    "vu_device_check": [],
    "vu_conv2d_fwd": [C.POINTER(ConvArgs), p],
    "vu_border_bias": [p, i, i, i, i, p, i, p, p],
    "vu_border_wsum": [p, i, i, i, i, p, p],
    "vu_border_bias_apply": [p, p, i, i, i, p, i, p],
    "vu_conv2d_wgrad": [C.POINTER(ConvArgs), p, i, i, p, p],
    "vu_conv2d_wgrad_oihw": [C.POINTER(ConvArgs), p, i, i, p, p],
    "vu_pack_weights_batched": [p, p, p, i, i, p],
    "vu_unpack_wgrad_batched": [p, p, p, i, i, p],
    "vu_pack_dgrad_weight": [p, p, i, i, i, i, i, p],
    "vu_pack_conv_weight": [p, p, i, i, i, i, i, i, p],
    "vu_pack_convT2x2_weight": [p, p, i, i, i, p],
    "vu_nchw_to_nhwc": [p, p, i, i, i, i, i, i, i, p],
    "vu_nhwc_to_nchw": [p, p, i, i, i, i, i, i, i, p],
    "vu_stem_conv7x7s2": [p, p, p, p, p, i, i, i, i, i, i, i, i, p],
    "vu_stem_conv7x7s2_act": [p, p, p, p, p, i, i, i, i, i, i, i, i, i, p],
    "vu_conv3x3_small_cin": [p, p, p, p, p, i, i, i, i, i, i, i, i, i, p],
    "vu_maxpool_nhwc": [p, p, i, i, i, i, i, i, i, i, i, i, i, i, p],
    "vu_bilinear_ac_nhwc": [p, p, i, i, i, i, i, i, i, i, i, i, i, i, i, i, i, p],
    "vu_broadcast_vec_nhwc": [p, p, i, i, i, i, i, i, i, p],
    "vu_attention_gate": [p, i, i, p, i, i, p, i, i, p, f, f, p, i, i, p, i, i, i, i, i, i, i, p],
    "vu_decoder_concat": [p, i, i, i, i, i, p, i, i, p, i, i, p, i, i, i, i, p, f, f, p, i, p, i, p, i, i, i, i, i, p],
    "vu_latent_head": [p, i, i, i, i, i, i, p, p, p, p, p, f, i, i, p, p, p, i, p],
    "vu_reparameterize": [p, p, p, f, p, sz, i, p],
    "vu_latent_project": [p, p, p, p, p, i, i, i, i, p],
    "vu_final_conv1": [p, i, i, i, i, i, i, p, f, p, p, i, p],
    "vu_resize_logits": [p, p, i, i, i, i, i, i, p],
    "vu_uncertainty_reduce": [p, i, i, i, i, i, i, i, p, p, p, p, p, p, p],
    "vu_moments_partial": [p, i, sz, p, p, p, p],
    "vu_moments_finalize": [p, p, p, p, i, sz, p, p, p, p, p, p],
    "vu_blend_accumulate": [p, i, i, i, i, i, i, i, i, i, p, p, i, i, p],
    "vu_blend_finalize": [p, p, p, sz, p],
    "vu_blend_accumulate_batch": [p, p, i, i, i, i, i, i, i, p, p, i, i, i, i, i, i, p],
    "vu_blend_finalize_shared": [p, p, p, sz, i, p],
    "vu_resample_u8": [p, i, i, i, p, i, i, i, p, p, i, p],
    "vu_nearest_u8": [p, i, i, i, p, i, i, p, p, p],
    "vu_tile_gather": [p, i, i, i, i, p, i, i, i, f, i, p, p],
    "vu_tile_stats": [p, i, i, i, i, p, f, p, p],
    "vu_seg_loss_sums": [p, p, sz, i, f, f, p, p],
    "vu_seg_loss_bwd": [p, p, sz, p, f, f, f, f, f, f, f, p, p],
    "vu_kl_free_bits": [p, p, i, i, f, p, p, p, f, p],
    "vu_ece_hist": [p, p, sz, i, p, p],
    "vu_confusion_counts": [p, p, sz, f, p, p],
    # ---- sort-based uncertainty metrics
    "vu_sort_pairs_desc_f32": [p, p, sz, p, p, p, sz, p],
    "vu_sort_pairs_desc_f32_static": [p, p, sz, p, p, p],
    "vu_pixel_nll_errors": [p, p, sz, f, p, p, p],
    "vu_sorted_bin_sums": [p, sz, p, i, p, p],
    "vu_masked_sum": [p, p, sz, p, p],
    "vu_error_labels": [p, p, sz, f, p, p],
    "vu_binary_curve_auc": [p, p, sz, p, sz, p, p],
    # ---- training step
    "vu_bn_stats": [p, sz, i, i, i, p, i, p],
    "vu_bn_finalize": [p, C.c_double, i, f, f, p, p, p, p, p],
    "vu_bn_apply": [p, sz, i, i, i, p, p, p, p, p, i, i, i, p, i, i, i, p],
    "vu_bn_update_running": [p, p, C.c_double, i, f, f, p, p, p],
    "vu_bn_apply_sums": [p, sz, i, i, i, p, f, f, p, p, p, p, p, p, p, i, i, i, p, i, i, i, p],
    "vu_bn_bwd": [p, sz, i, i, i, p, i, i, p, i, i, p, p, p, p, p, i, i, p, i, i, p, p, p, i, p],
    "vu_add_nhwc": [p, i, i, p, i, i, sz, i, i, p],
    "vu_maxpool_bwd": [p, p, p, i, i, i, i, i, i, i, i, i, i, i, i, i, i, p],
    "vu_maxpool_idx_fwd": [p, p, p, i, i, i, i, i, i, i, i, i, i, i, i, p],
    "vu_maxpool_idx_bwd": [p, p, p, i, i, i, i, i, i, i, i, i, i, i, i, p],
    "vu_bilinear_ac_bwd": [p, p, i, i, i, i, i, i, i, i, i, i, i, i, p],
    "vu_resize_map_bwd": [p, p, i, i, i, i, i, p],
    "vu_scale_by_map": [p, i, i, p, p, i, i, sz, i, i, i, p],
    "vu_channel_dot": [p, i, i, p, i, i, p, sz, i, i, p],
    "vu_map_expand": [p, p, p, i, i, p, i, i, sz, i, i, p],
    "vu_weighted_channel_sum": [p, i, i, p, i, sz, i, i, p, i, p],
    "vu_map_bn_sigmoid_fwd": [p, sz, p, p, f, f, p, p, p, p, p, p],
    "vu_map_bn_sigmoid_bwd": [p, p, p, p, p, sz, p, p, p, p, p],
    "vu_map_sum": [p, sz, p, p],
    "vu_vec_layer_fwd": [p, p, p, i, i, i, p, p, f, f, C.c_double, p, p, i, p, p, p, p, p],
    "vu_vec_layer_bwd": [p, p, p, p, p, i, i, i, p, p, p, i, p, p, p, p, p, p, p],
    "vu_reparam_bwd": [p, p, p, f, p, p, sz, i, p],
    "vu_im2col_small_cin": [p, p, i, i, i, i, i, i, i, i, i, p],
    "vu_pack_rows_padded": [p, p, i, i, i, i, p],
    "vu_stem_wgrad": [p, p, i, i, p, i, i, i, i, i, i, p],
    "vu_sumsq": [p, sz, p, i, p],
    "vu_adamw_step": [p, p, p, p, sz, f, f, f, f, f, i, p, f, f, p],
}

_lib = None


def load() -> C.CDLL:
    """Load the extension; fail loudly if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise VaeUnetError(
            f"{LIB_PATH} is not built. Run `python -m vaeunet_b200.build` (needs nvcc). "
            "vaeunet_b200 has no CPU / eager fallback.")
    lib = C.CDLL(str(LIB_PATH))
    lib.vu_last_error.restype = C.c_char_p
    lib.vu_last_error.argtypes = []
    lib.vu_sort_workspace_bytes.restype = C.c_size_t
    lib.vu_sort_workspace_bytes.argtypes = [C.c_size_t]
    for name, argt in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so is stale
        fn.restype = C.c_int
        fn.argtypes = argt
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().vu_last_error().decode("utf-8", "replace")
        raise VaeUnetError(f"{what} failed (vu_status {rc}): {msg}")


_device_ok = None


def require_device() -> None:
    """Raise unless a B200-class device is current (product path has no fallback)."""
    global _device_ok
    if _device_ok:
        return
    import torch
    if not torch.cuda.is_available():
        raise VaeUnetError("vaeunet_b200 needs a CUDA sm_100 (B200) device; none is visible and there is no CPU fallback")
    check(load().vu_device_check(), "vu_device_check")
    _device_ok = True
