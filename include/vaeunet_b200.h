/* vaeunet_b200 — C ABI of the B200-native VAE-UNet hot path.
 *
 * The reference (tmuird/VAEUNET) has no FFI layer: its boundary is the PyTorch module surface
 * (SURVEY.md 8b). This header is what the torch custom-op shims in vaeunet_b200/ops.py bind
 * (ctypes); every entry point names the reference operation it replaces (file:line in the
 * upstream repo). Conventions:
 *   - all pointers are DEVICE pointers unless the name ends in _host; activations are
 *     NHWC-contiguous 16-bit (bf16 by default, fp16 when dtype == VU_F16) unless stated;
 *   - `ctot` is the channel pitch of a buffer, `coff` the first channel of the slice used
 *     (concat is done by writing into channel slices; torch.cat never runs);
 *   - kernels are enqueued on `stream` and never synchronise; no allocation happens inside;
 *   - return 0 on success or a negative vu_status; vu_last_error() gives a thread-local text.
 */
#ifndef VAEUNET_B200_H
#define VAEUNET_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* vu_stream_t; /* cudaStream_t */

enum vu_status {
    VU_OK = 0,
    VU_ERR_INVALID = -1,     /* bad shape / argument */
    VU_ERR_ALIGN = -2,       /* misaligned pointer or channel offset */
    VU_ERR_UNSUPPORTED = -3, /* combination not implemented */
    VU_ERR_CUDA = -4,        /* CUDA runtime / driver error */
    VU_ERR_NO_DEVICE = -5    /* no sm_100 device: there is no CPU fallback */
};
enum vu_dtype { VU_BF16 = 0, VU_F16 = 1 };

const char* vu_last_error(void);
int vu_version(void);
/* 0 when the current device is sm_100 (B200), VU_ERR_NO_DEVICE otherwise. */
int vu_device_check(void);

/* ---------------------------------------------------------------- dense contractions
 * Implicit-GEMM convolution on tcgen05/TMEM, operands staged by TMA (128B swizzle).
 * Replaces nn.Conv2d (+ folded eval BatchNorm2d + ReLU + residual add) everywhere on the path:
 *   3x3 s1 p1  unet/unet_parts.py:40,43 ; unet/unet_resnet.py:60,66 ; ResNet BasicBlock convs
 *   3x3 s2 p1 / 1x1 s2 ResNet stage entry + downsample shortcut (timm call site unet_resnet.py:131)
 *   1x1 s1     attention W_g / W_x unet/unet_parts.py:10-17, unet/unet_resnet.py:9-16
 *   ConvTranspose2d k2 s2 (unet/unet_parts.py:76) as four 1x1 GEMMs with a pixel-shuffle store
 *     (out_up = 2, out_dh/out_dw select the sub-pixel).
 * y[n, ho*out_up+out_dh, wo*out_up+out_dw, y_coff+co] =
 *     act( scale[co] * sum_{r,s,ci} x[n, ho*stride+r-pad, wo*stride+s-pad, x_coff+ci] * w[co][r*KW+s][ci]
 *          + bias[co] + residual[n,ho,wo,res_coff+co] )
 */
typedef struct vu_conv_args {
    const void* x; int N, H, W, x_ctot, x_coff, Cin;
    const void* w;            /* packed [Cout][KH*KW][Cin], 16-bit (vu_pack_conv_weight) */
    int Cout, KH, KW, stride, pad;
    const float* scale;       /* [Cout] fp32 or NULL (1)  — folded BN gamma/sqrt(var+eps) */
    const float* bias;        /* [Cout] fp32 or NULL (0)  — folded BN beta - mean*scale (+ conv bias) */
    const void* residual; int res_ctot, res_coff;  /* NHWC at output resolution, or NULL */
    int relu;
    void* y; int y_ctot, y_coff;
    int out_up, out_dh, out_dw;  /* 1,0,0 for ordinary convs */
    int dtype;                   /* vu_dtype */
    int max_ctas;                /* 0 = one persistent CTA per SM */
    /* Optional fused output conv (final_conv unet_resnet.py:189,237 / OutConv unet_parts.py:97-103, n_classes = 1):
     * fuse_out[n,ho,wo] = fuse_bias + sum_co fuse_w[co] * act(...)[co] in fp32, taken before the 16-bit rounding of
     * y. Only for 3x3 s1 p1 convs with Cout in {64,128} on maps at least 96 wide (the smem tap-reuse kernel);
     * y may then be NULL (activation not materialised). NULL = off. */
    const float* fuse_w; float fuse_bias; float* fuse_out;
    /* Optional position-dependent pre-scale bias, fp32 [N][9][Cout] (3x3 s1 p1 convs only): the exact contribution of
     * input channels that are spatially CONSTANT per image (z_proj(z) in a DecoderBlock's concat, unet_resnet.py:93-98)
     * and were therefore left out of x / w. Class = 3*rowclass + colclass of the output pixel (0 first row/col, 2 last,
     * 1 interior): zero padding removes different taps on each border (vu_border_bias builds the table). NULL = off.
     *   y = act(scale * (conv(x, w) + pos_bias[n][class(ho,wo)][co]) + bias + residual) */
    const float* pos_bias;
    /* Optional per-channel statistics of the STORED output (after rounding to 16 bit): stat_sums[co] += sum y,
     * stat_sums[Cout + co] += sum y^2 over all output pixels, fp64, accumulated (zero it first). The training step feeds them
     * to vu_bn_finalize, so train-mode BatchNorm needs no separate statistics pass over the conv output. NULL = off. */
    double* stat_sums;
    /* Optional bit mask of the filter taps to run (bit r*KW+s; 0 = all taps): taps whose filter rows are identically zero are
     * skipped. The four parity banks of a stride-2 data gradient (vu_pack_dgrad_weight) hold 1, 2, 2 and 4 non-zero taps of 9. */
    int tap_mask;
} vu_conv_args;
int vu_conv2d_fwd(const vu_conv_args* a, vu_stream_t stream);

/* pos_bias table for vu_conv_args: out[d][cls][co] = sum over the taps (r,s) that stay inside the image for border class
 * cls, and over c < Cz, of w[co][coff + c][r][s] * z[d][c]. w: the conv's OIHW fp32 weight [Cout][Cin_total][3][3];
 * z fp32 [D][Cz] (the constant channel values). */
int vu_border_bias(const float* w_oihw, int Cout, int Cin_total, int coff, int Cz, const float* z, int D, float* out,
                   vu_stream_t stream);

/* The same table in two steps for fixed weights (inference): wsum fp32 [9][Cout][Cz] once, then one dot product per
 * (image, class, output channel) per call (Cz % 4 == 0). */
int vu_border_wsum(const float* w_oihw, int Cout, int Cin_total, int coff, int Cz, float* wsum, vu_stream_t stream);
int vu_border_bias_apply(const float* wsum, const float* z, int D, int Cout, int Cz, float* out, int accumulate /* out += */,
                         vu_stream_t stream);

/* Weight gradient of the same convolution (training step, train.py:397-399 loss.backward()):
 *   dw[co][r*KW+s][ci] += sum_{n,ho,wo} dy[n,ho,wo,dy_coff+co] * x[n, ho*stride+r-pad, wo*stride+s-pad, x_coff+ci]
 * `a` gives x and the geometry (x, N,H,W,x_ctot,x_coff,Cin,Cout,KH,KW,stride,pad,dtype; other fields ignored);
 * dy is the 16-bit NHWC gradient of the conv output [N,Ho,Wo,dy_ctot]. dw is fp32 in the PACKED order
 * [Cout][KH*KW][Cin] and is ACCUMULATED into (K splits combine with atomics): zero it first. */
int vu_conv2d_wgrad(const vu_conv_args* a, const void* dy, int dy_ctot, int dy_coff, float* dw, vu_stream_t stream);
/* Same, accumulated straight into the parameter's own OIHW layout dw[co][ci][r][s] (the .grad of nn.Conv2d.weight). */
int vu_conv2d_wgrad_oihw(const vu_conv_args* a, const void* dy, int dy_ctot, int dy_coff, float* dw, vu_stream_t stream);
/* vu_pack_conv_weight + vu_pack_dgrad_weight for a whole list of convolutions in one launch (the training step repacks
 * every bank after each optimizer step). table: device array of records {int64 src_off (floats from src_base),
 * int64 dst_fwd_off, int64 dst_bank_off (16-bit elements from dst_base), int32 O, I, K, stride}. */
int vu_pack_weights_batched(const float* src_base, void* dst_base, const void* table, int n_entries, int dtype,
                            vu_stream_t stream);
/* Packed weight gradients (vu_conv2d_wgrad) of a list of convolutions -> the parameters' OIHW gradient buffers, accumulated,
 * in one launch. table: device array of n_entries records {int64 src_off, int64 dst_off (floats from the two bases),
 * int32 O, T, I, pad}; max_row_floats = max over entries of T*I. */
int vu_unpack_wgrad_batched(const float* packed_base, float* oihw_base, const void* table, int n_entries,
                            int max_row_floats, vu_stream_t stream);
/* Filter banks for the DATA gradient of conv2d(x, w) (w OIHW fp32; 3x3 p1 or 1x1 p0; stride 1|2), to be run by
 * vu_conv2d_fwd on dY: stride 1 -> one bank [I][K*K][O] (taps rotated 180 degrees); stride 2, 1x1 -> one bank [I][1][O]
 * (written to dX parity (0,0) with out_up = 2); stride 2, 3x3 -> four 3x3 p1 banks [ph*2+pc][I][9][O], one per parity of
 * the dX pixel (out_up = 2, out_dh = ph, out_dw = pc). */
int vu_pack_dgrad_weight(const float* w_oihw, void* banks, int O, int I, int K, int stride, int dtype, vu_stream_t stream);

/* OIHW fp32 -> [O][KH*KW][I] 16-bit. transpose_io: weight is IOHW (ConvTranspose2d). */
int vu_pack_conv_weight(const float* w_oihw, void* w_packed, int O, int I, int KH, int KW,
                        int transpose_io, int dtype, vu_stream_t stream);
/* ConvTranspose2d k2 s2 weight [I][O][2][2] fp32 -> four packed 1x1 filters [4][O][1][I]. */
int vu_pack_convT2x2_weight(const float* w_iohw, void* w_packed, int I, int O, int dtype,
                            vu_stream_t stream);

/* ---------------------------------------------------------------- layout / pooling / resize */
/* NCHW fp32 -> NHWC 16-bit into a channel slice (first op on images; train.py:382). */
int vu_nchw_to_nhwc(const float* x, void* y, int N, int C, int H, int W, int y_ctot, int y_coff,
                    int dtype, vu_stream_t stream);
/* NHWC 16-bit slice -> NCHW fp32 (module-surface returns). */
int vu_nhwc_to_nchw(const void* x, float* y, int N, int C, int H, int W, int x_ctot, int x_coff,
                    int dtype, vu_stream_t stream);
/* ResNet stem conv 7x7 s2 p3 (Cin <= 4) + folded BN + ReLU, NCHW fp32 in -> NHWC 16-bit out.
 * (timm resnet34 conv1/bn1/act1 reached from unet/unet_resnet.py:131-137,201). w: OIHW fp32. */
int vu_stem_conv7x7s2(const float* x_nchw, const float* w_oihw, const float* scale, const float* bias,
                      void* y, int N, int Cin, int H, int W, int Cout, int y_ctot, int y_coff,
                      int dtype, vu_stream_t stream);
/* Same with the activation selectable (relu = 0: the raw conv output that train-mode BatchNorm consumes). */
int vu_stem_conv7x7s2_act(const float* x_nchw, const float* w_oihw, const float* scale, const float* bias,
                          void* y, int N, int Cin, int H, int W, int Cout, int y_ctot, int y_coff, int relu,
                          int dtype, vu_stream_t stream);
/* Direct 3x3 s1 p1 conv for tiny Cin (basic UNet `inc`, unet/unet_parts.py:40 with Cin = n_channels). */
int vu_conv3x3_small_cin(const float* x_nchw, const float* w_oihw, const float* scale, const float* bias,
                         void* y, int N, int Cin, int H, int W, int Cout, int y_ctot, int y_coff,
                         int relu, int dtype, vu_stream_t stream);
/* MaxPool2d(k, s, p) on NHWC 16-bit: (3,2,1) ResNet stem pool; (2,2,0) unet/unet_parts.py:57. */
int vu_maxpool_nhwc(const void* x, void* y, int N, int H, int W, int C, int x_ctot, int x_coff,
                    int k, int s, int p, int y_ctot, int y_coff, int dtype, vu_stream_t stream);
/* F.interpolate(mode='bilinear', align_corners=True) NHWC 16-bit -> channel slice
 * (unet/unet_resnet.py:79; nn.Upsample unet/unet_parts.py:73). pad_*: F.pad offsets of Up
 * (unet/unet_parts.py:88) — destination is [N,Hd,Wd], the resized map lands at (pad_t,pad_l). */
int vu_bilinear_ac_nhwc(const void* x, void* y, int N, int Hs, int Ws, int C, int x_ctot, int x_coff,
                        int Ho, int Wo, int Hd, int Wd, int pad_t, int pad_l, int y_ctot, int y_coff,
                        int dtype, vu_stream_t stream);
/* Broadcast per-image vectors v[Nimg][C] (fp32) over H x W into a channel slice: the spatially
 * constant maps z_initial(z_spatial) and z_proj(z) (unet/unet_resnet.py:93-94,221-226). */
int vu_broadcast_vec_nhwc(const float* v, void* y, int N, int H, int W, int C, int y_ctot, int y_coff,
                          int dtype, vu_stream_t stream);

/* ---------------------------------------------------------------- attention gate
 * out = skip * sigmoid(psi_scale * <psi_w, relu(g1 + x1)> + psi_bias)
 * (unet/unet_parts.py:24-30 == unet/unet_resnet.py:24-29). g1 = BN(W_g g) [Ng,H,W,F] and
 * x1 = BN(W_x skip) [Ns,H,W,F] come from vu_conv2d_fwd; Ng = Ns * rep (rep MC samples share one
 * skip/x1). psi_w fp32 [F]; psi (optional, fp32 [Ng,H,W]) keeps the gate map for hooks. */
int vu_attention_gate(const void* g1, int g1_ctot, int g1_coff, const void* x1, int x1_ctot, int x1_coff,
                      const void* skip, int skip_ctot, int skip_coff, const float* psi_w,
                      float psi_scale, float psi_bias, void* out, int out_ctot, int out_coff,
                      float* psi, int Ng, int rep, int H, int W, int F, int Cl, int dtype,
                      vu_stream_t stream);

/* The whole concat buffer of a DecoderBlock in one pass (unet/unet_resnet.py:71-98):
 *   out[d,ho,wo,:] = [ bilinear_ac(x_low) | skip * sigmoid(psi_scale*<psi_w, relu(bilinear_ac(g1_low)+x1)> + psi_bias) | zvec[d,:] ]
 * g1_low = BN(W_g x_low) is taken at the LOW resolution of x_low (a 1x1 conv + affine commutes with bilinear
 * interpolation, whose weights sum to 1) and interpolated on the fly; a spatially constant x is a 1x1 map.
 * x_low [D,hs,ws,Cx], g1_low [D,hs,ws,F], x1 [D/rep,H,W,F], skip [D/rep,H,W,Cs] 16-bit NHWC slices; zvec fp32 [D,Cz]
 * or NULL (Cz = 0); out [D,H,W,out_ctot] receives channels [0, Cx+Cs+Cz); psi_out fp32 [D,H,W] or NULL. */
int vu_decoder_concat(const void* x_low, int hs, int ws, int Cx, int x_ctot, int x_coff, const void* g1_low,
                      int g1_ctot, int g1_coff, const void* x1, int x1_ctot, int x1_coff, const void* skip,
                      int skip_ctot, int skip_coff, int F, int Cs, const float* psi_w, float psi_scale,
                      float psi_bias, const float* zvec, int Cz, void* out, int out_ctot, float* psi_out,
                      int D, int rep, int H, int W, int dtype, vu_stream_t stream);

/* ---------------------------------------------------------------- latent head
 * mu/logvar = GAP(conv1x1(f4) + b) = W * mean_hw(f4) + b (unet/unet_resnet.py:140-147,205-206;
 * utils/vae_utils.py:13-25), then for every MC sample n: z[n,b,:] = mu + eps[n,b,:]*exp(.5 logvar)*T
 * (utils/vae_utils.py:5-10,47-50; unet_resnet.py:191-194). f4 NHWC 16-bit [B,h,w,C]; w_mu/w_lv fp32
 * [L][C]; eps fp32 [Ns,B,L] or NULL (z = mu). Outputs fp32: mu,logvar [B,L]; z [Ns,B,L]. */
int vu_latent_head(const void* f4, int B, int h, int w, int C, int f_ctot, int f_coff,
                   const float* w_mu, const float* b_mu, const float* w_lv, const float* b_lv,
                   const float* eps, float temperature, int Ns, int L,
                   float* mu, float* logvar, float* z, int dtype, vu_stream_t stream);
/* z[n, j] = mu[j] + eps[n, j] * exp(.5 logvar[j]) * T  (utils/vae_utils.py:5-10; unet_resnet.py:191-194).
 * mu/logvar fp32 [n_per_sample]; eps/z fp32 [num_samples][n_per_sample]. */
int vu_reparameterize(const float* mu, const float* logvar, const float* eps, float temperature, float* z,
                      size_t n_per_sample, int num_samples, vu_stream_t stream);
/* v[i,:] = act(scale * (W z[i,:] + b) + shift): the 1x1 conv + BN + ReLU applied to a spatially
 * constant map (z_initial unet_resnet.py:150-154; z_proj :37-41). z fp32 [M][L], W fp32 [C][L]. */
int vu_latent_project(const float* z, const float* w, const float* scale, const float* shift,
                      float* v, int M, int L, int C, int relu, vu_stream_t stream);

/* ---------------------------------------------------------------- output head + uncertainty
 * final_conv 1x1 C->1 + bias on NHWC 16-bit -> fp32 logits [N,H,W] (unet_resnet.py:189,237;
 * OutConv unet/unet_parts.py:97-103 for n_classes = 1). bias_dev (optional, device fp32[1]) is added to `bias`:
 * the training step reads the live parameter without a host round trip. */
int vu_final_conv1(const void* x, int N, int H, int W, int C, int x_ctot, int x_coff,
                   const float* w, float bias, const float* bias_dev, float* logits, int dtype, vu_stream_t stream);
/* bilinear (align_corners=True) resize of fp32 maps [N,Hs,Ws] -> [N,Ho,Wo], optional sigmoid
 * before (predict_with_patches order, visualize_vae.py:345-358) or after (forward /
 * predict_full_image order, unet_resnet.py:238; visualize_vae.py:81-83). mode: 0 none, 1 pre, 2 post */
int vu_resize_logits(const float* x, float* y, int N, int Hs, int Ws, int Ho, int Wo, int sigmoid_mode,
                     vu_stream_t stream);
/* calculate_uncertainty_metrics (visualize_vae.py:90-117) fused with the final upsample+sigmoid:
 * logits fp32 [B, Ns, Hs, Ws] (sample-minor per patch) -> per-pixel over Ns:
 * mean, unbiased std, entropy of mean, mutual information, coeff. of variation, each [B,Ho,Wo] fp32
 * (any output pointer may be NULL). input_is_prob: skip resize+sigmoid (maps already [B,Ns,Ho,Wo]).
 * probs_out (optional) receives the per-sample probability maps [B,Ns,Ho,Wo]. */
int vu_uncertainty_reduce(const float* logits, int B, int Ns, int Hs, int Ws, int Ho, int Wo,
                          int input_is_prob, float* mean, float* stdev, float* entropy,
                          float* mutual_info, float* coeff_var, float* probs_out, vu_stream_t stream);
/* Chan merge of per-shard moments for sample-sharded MC: (count, mean, M2, sumH). */
int vu_moments_partial(const float* probs, int Ns, size_t npix, float* mean, float* m2, float* sum_ent,
                       vu_stream_t stream);
int vu_moments_finalize(const float* counts_host, const float* mean, const float* m2, const float* sum_ent,
                        int nshards, size_t npix, float* out_mean, float* out_std, float* out_entropy,
                        float* out_mi, float* out_cv, vu_stream_t stream);

/* ---------------------------------------------------------------- tile blending
 * predict_with_patches accumulation (visualize_vae.py:361-384,409): ramp weights are analytic in
 * tile coordinates. acc/wsum fp32 [H,W]; tile fp32 [ph,pw] probabilities. */
int vu_blend_accumulate(const float* tile, int ph, int pw, int sh, int sw, int ti, int tj, int nth, int ntw,
                        int overlap, float* acc, float* wsum, int H, int W, vu_stream_t stream);
int vu_blend_finalize(const float* acc, const float* wsum, float* out, size_t n, vu_stream_t stream);

/* One launch for a batch of T equal-sized tiles x Ns MC samples (tiles fp32 [T][Ns][ph][pw]; info int32 [T][4] =
 * {sh, sw, ti, tj} on the device): every pixel of the batch's bounding box [y0,y1) x [x0,x1) is owned by one thread that walks
 * the tiles in index order (no atomics; the same fp32 summation order as one vu_blend_accumulate per tile). acc fp32
 * [Ns][H][W]; wsum fp32 [H][W] - ONE weight map, the ramp weights do not depend on the sample. */
int vu_blend_accumulate_batch(const float* tiles, const int* info, int T, int Ns, int ph, int pw, int nth, int ntw,
                              int overlap, float* acc, float* wsum, int H, int W, int y0, int y1, int x0, int x1,
                              vu_stream_t stream);
/* out[n][i] = acc[n][i] / (wsum[i] + 1e-8), i < plane (visualize_vae.py:409 with the shared weight map). */
int vu_blend_finalize_shared(const float* acc, const float* wsum, float* out, size_t plane, int Ns, vu_stream_t stream);

/* ---------------------------------------------------------------- patch pipeline
 * Device side of utils/data_loading.py:302-446,580-601 (resize -> /255 -> 50 %-overlap tiling -> black-border / lesion
 * flags) and of the tile cutting of predict_with_patches (visualize_vae.py:253-300): the decoded 8-bit image goes to the
 * device once, everything after it stays there.
 * vu_resample_u8: ONE pass (axis 0 = along x, 1 = along y) of Pillow's 8-bit ImagingResample on an HWC image with
 * host-built tables: bounds int32 [out][2] = {first input index, taps}, kk int32 [out][ksize] = coefficients in 22-bit
 * fixed point; the result is rounded and clipped to 8 bits like Pillow's intermediate and final images (bit-exact with
 * PIL.Image.resize for BICUBIC / BILINEAR / LANCZOS / BOX when the tables come from the same formulas).
 * vu_nearest_u8: dst[y][x] = src[yi[y]][xi[x]] (Image.NEAREST with host-built index tables; masks). */
int vu_resample_u8(const uint8_t* src, int Hs, int Ws, int C, uint8_t* dst, int Hd, int Wd, int axis, const int* bounds,
                   const int* kk, int ksize, vu_stream_t stream);
int vu_nearest_u8(const uint8_t* src, int Hs, int Ws, int C, uint8_t* dst, int Hd, int Wd, const int* yi, const int* xi,
                  vu_stream_t stream);
/* out fp32 NCHW [T][C][ph][pw] = tiles of one image at coords int32 [T][2] = {y0, x0} (device). src: uint8 HWC [H][W][C]
 * (src_is_u8; values are divided by `divisor`, 255 for images) or fp32 CHW [C][H][W]. binarise: out = (value > 0) ? 1 : 0
 * (the reference's mask conversion). Pixels outside the image are 0. */
int vu_tile_gather(const void* src, int src_is_u8, int C, int H, int W, const int* coords, int T, int ph, int pw,
                   float divisor, int binarise, float* out, vu_stream_t stream);
/* stats fp32 [T][2]: {fraction of pixels whose channel mean < black_thr (is_valid_patch, data_loading.py:287-300), 1 if any
 * mask pixel > 0.5 (has_lesion, :382) else 0}. masks fp32 [T][ph][pw] or NULL. */
int vu_tile_stats(const float* tiles, int T, int C, int ph, int pw, const float* masks, float black_thr, float* stats,
                  vu_stream_t stream);

/* ---------------------------------------------------------------- losses / metrics
 * One pass over logits+targets -> 8 partial sums (fp64 accumulators in `sums`):
 *   [0] sum p, [1] sum t, [2] sum p*t, [3] sum BCE-with-logits, [4] sum focal term,
 *   [5] count, [6] sum (p-t)^2 (Brier), [7] sum NLL(clamp(p,1e-9))
 * (utils/loss.py:6-28,54-63,75-92; utils/uncertainty_metrics.py:137-139,334-338).
 * input_is_prob: x already holds probabilities (metrics on mean maps). */
int vu_seg_loss_sums(const float* x, const float* target, size_t n, int input_is_prob,
                     float focal_alpha, float focal_gamma, double* sums, vu_stream_t stream);
/* d(loss)/d(logits) for  w_bce*BCE + w_dice*dice + w_focal*focal, from the sums above. */
int vu_seg_loss_bwd(const float* logits, const float* target, size_t n, const double* sums,
                    float w_bce, float w_dice, float w_focal, float focal_alpha, float focal_gamma,
                    float smooth, float grad_scale, float* dlogits, vu_stream_t stream);
/* kl_with_free_bits (utils/loss.py:148-170) forward value + analytic grads. */
int vu_kl_free_bits(const float* mu, const float* logvar, int B, int L, float free_bits,
                    float* kl_out, float* dmu, float* dlogvar, float grad_scale, vu_stream_t stream);
/* ECE histogram (utils/uncertainty_metrics.py:108-135): per-bin count, sum gt, sum p (fp64 [3*bins]). */
int vu_ece_hist(const float* p, const float* gt, size_t n, int num_bins, double* hist, vu_stream_t stream);
/* Thresholded confusion counts (utils/metrics.py:98-117; uncertainty_metrics.py:340-353):
 * counts[0..3] = tp, fp, tn, fn of (x > thr) vs (t > 0.5) as fp64. */
int vu_confusion_counts(const float* x, const float* t, size_t n, float thr, double* counts,
                        vu_stream_t stream);

/* ---------------------------------------------------------------- sort-based uncertainty metrics
 * Device replacements for numpy argsort + sklearn in utils/uncertainty_metrics.py:175-240 (sparsification curve) and
 * :270-283 (AUROC / AUPRC of "uncertainty predicts error").
 * vu_sort_pairs_desc_f32: stable LSD radix sort of (score, 32-bit payload) pairs by score DESCENDING; keys_sorted receives
 * order-preserving unsigned keys (equal keys <=> equal scores). workspace: vu_sort_workspace_bytes(n) bytes, shared with
 * vu_binary_curve_auc. */
size_t vu_sort_workspace_bytes(size_t n);
int vu_sort_pairs_desc_f32(const float* score, const uint32_t* vals, size_t n, uint32_t* keys_sorted, uint32_t* vals_sorted,
                           void* workspace, size_t workspace_bytes, vu_stream_t stream);
/* vu_sort_pairs_desc_f32_static: the same sort with NO caller workspace and no allocation of any kind - its scratch is a
 * file-scope device array inside the library. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid allocation).
 * Limit: n <= VU_SORT_STATIC_MAX_N pairs (one 512 x 512 map; larger n is rejected with VU_ERR_UNSUPPORTED); the 6.1 MB array is
 * held by every GPU that loads the library until the process exits. All calls share it: calls that overlap on the GPU
 * (different streams / threads) race - serialise them on one stream. */
#define VU_SORT_STATIC_MAX_N 262144
int vu_sort_pairs_desc_f32_static(const float* score, const uint32_t* vals, size_t n, uint32_t* keys_sorted,
                                  uint32_t* vals_sorted, vu_stream_t stream);
/* err[i] = -(g log(p + eps) + (1 - g) log(1 - p + eps)) (uncertainty_metrics.py:179-181); stats: 4 doubles -
 * [0] sum of the non-NaN errors, [1] number of NaNs, and two uint32 in the low words of [2]: order-preserving encodings
 * (u ^ (sign ? ~0 : 1 << 31)) of the minimum and maximum error. */
int vu_pixel_nll_errors(const float* p, const float* gt, size_t n, float eps, float* err, double* stats, vu_stream_t stream);
/* sums[b] = sum of the float values vals_sorted[i], cuts[b] <= i < cuts[b+1] (device int64 cuts[nbins + 1], ascending). */
int vu_sorted_bin_sums(const uint32_t* vals_sorted, size_t n, const long long* cuts, int nbins, double* sums,
                       vu_stream_t stream);
/* out[0] = sum of x[i] over the set bits of a packed keep-mask (bit i & 7 of byte i >> 3): the random-removal baseline of the
 * sparsification curve, whose subsets come from numpy's RNG stream by the reference's definition (np.random.seed(i + 1000 b)). */
int vu_masked_sum(const float* x, const uint8_t* bits, size_t n, double* out, vu_stream_t stream);
/* labels[i] = ((p[i] > thr) != (gt[i] > 0.5)) (uncertainty_metrics.py:271-275). */
int vu_error_labels(const float* p, const float* gt, size_t n, float thr, uint32_t* labels, vu_stream_t stream);
/* Binary-classification curve of pairs sorted by score descending (sklearn _binary_clf_curve: cumulative TP / FP at the last
 * element of every group of tied scores): out[0] = un-normalised ROC trapezoid area, out[1] = un-normalised average
 * precision, out[2] = positives, out[3] = negatives; AUROC = out[0] / (out[2] out[3]), AP = out[1] / out[2]. */
int vu_binary_curve_auc(const uint32_t* keys_sorted, const uint32_t* labels_sorted, size_t n, void* workspace,
                        size_t workspace_bytes, double* out, vu_stream_t stream);

/* ---------------------------------------------------------------- training step (train.py:385-411)
 * Train-mode BatchNorm2d over an NHWC 16-bit slice (unet_parts.py:41,44; unet_resnet.py:61,67 and every timm BN):
 * vu_bn_stats -> sums[0..C) = sum x, sums[C..2C) = sum x^2 (fp64); vu_bn_finalize -> mean / invstd (biased var, eps)
 * and the running statistics update (momentum, unbiased var); vu_bn_apply -> y = act(gamma*(x-mean)*invstd + beta
 * + residual). vu_bn_bwd: dz = dy * (y_relu > 0 if y_relu), sums <- (sum dz, sum dz*xhat) = (dbeta, dgamma) in fp64,
 * dx = gamma*invstd*(dz - mean(dz) - xhat*mean(dz*xhat)); dz_out (optional) receives dz (residual-branch gradient).
 * relu_beta (optional, with y_relu NULL): the layer was BN + ReLU without a residual - the mask is recomputed from x with
 * the forward's own expressions (gamma*invstd*x + beta - mean*gamma*invstd > 0), saving the read of y in both passes. */
int vu_bn_stats(const void* x, size_t npix, int C, int x_ctot, int x_coff, double* sums, int dtype, vu_stream_t stream);
int vu_bn_finalize(const double* sums, double count, int C, float eps, float momentum, float* mean, float* invstd,
                   float* running_mean, float* running_var, vu_stream_t stream);
int vu_bn_apply(const void* x, size_t npix, int C, int x_ctot, int x_coff, const float* mean, const float* invstd,
                const float* gamma, const float* beta, const void* residual, int res_ctot, int res_coff, int relu,
                void* y, int y_ctot, int y_coff, int dtype, vu_stream_t stream);
/* running_mean / running_var update of nn.BatchNorm2d (momentum, unbiased variance) from saved batch statistics
 * (mean, invstd over `count` values per channel): the functional torch custom ops (vaeunet::batch_norm_train, ::vec_layer,
 * ::map_bn_sigmoid) leave the module buffers alone and vaeunet::bn_update_running mutates them. */
int vu_bn_update_running(const float* mean, const float* invstd, double count, int C, float eps, float momentum,
                         float* running_mean, float* running_var, vu_stream_t stream);
/* vu_bn_finalize + vu_bn_apply in ONE launch: every CTA derives mean / invstd of its channels from the fp64 sums (the
 * statistics pass or the producing conv's epilogue, vu_conv_args.stat_sums); count = npix. mean / invstd [C] are written
 * for the backward and the running statistics are updated (NULL = leave them). */
int vu_bn_apply_sums(const void* x, size_t npix, int C, int x_ctot, int x_coff, const double* sums, float eps, float momentum,
                     float* running_mean, float* running_var, float* mean, float* invstd, const float* gamma,
                     const float* beta, const void* residual, int res_ctot, int res_coff, int relu, void* y, int y_ctot,
                     int y_coff, int dtype, vu_stream_t stream);
int vu_bn_bwd(const void* x, size_t npix, int C, int x_ctot, int x_coff, const void* dy, int dy_ctot, int dy_coff,
              const void* y_relu, int y_ctot, int y_coff, const float* mean, const float* invstd, const float* gamma,
              double* sums, void* dx, int dx_ctot, int dx_coff, void* dz_out, int dz_ctot, int dz_coff,
              float* dgamma /* += , optional */, float* dbeta /* += */, const float* relu_beta, int dtype, vu_stream_t stream);
/* dst += src (gradient accumulation over the consumers of an activation). */
int vu_add_nhwc(void* dst, int d_ctot, int d_coff, const void* src, int s_ctot, int s_coff, size_t npix, int C,
                int dtype, vu_stream_t stream);
/* Backward of vu_maxpool_nhwc (first maximum wins, as ATen) and of vu_bilinear_ac_nhwc / vu_resize_logits (gathers). */
int vu_maxpool_bwd(const void* x, const void* dy, void* dx, int N, int H, int W, int C, int x_ctot, int x_coff,
                   int dy_ctot, int dy_coff, int dx_ctot, int dx_coff, int k, int s, int p, int dtype, vu_stream_t stream);
/* Max-pool that also saves, per output element, the window position r*k+q of its first maximum (uint8 [N,Ho,Wo,C]),
 * and the backward as a gather over those codes (training fast path; same results as vu_maxpool_nhwc / vu_maxpool_bwd). */
int vu_maxpool_idx_fwd(const void* x, void* y, void* code, int N, int H, int W, int C, int x_ctot, int x_coff,
                       int y_ctot, int y_coff, int k, int s, int p, int dtype, vu_stream_t stream);
int vu_maxpool_idx_bwd(const void* code, const void* dy, void* dx, int N, int H, int W, int C, int dy_ctot, int dy_coff,
                       int dx_ctot, int dx_coff, int k, int s, int p, int dtype, vu_stream_t stream);
int vu_bilinear_ac_bwd(const void* dy, void* dx, int N, int Hs, int Ws, int C, int Ho, int Wo, int dy_ctot, int dy_coff,
                       int dx_ctot, int dx_coff, int accumulate, int dtype, vu_stream_t stream);
int vu_resize_map_bwd(const float* dy, float* dx, int N, int Hs, int Ws, int Ho, int Wo, vu_stream_t stream);
/* Per-pixel pieces of the attention gate / output conv and their gradients:
 *   scale_by_map: out[p,c] (+)= a[p,c]*m[p];  channel_dot: out[p] = sum_c a[p,c]*b[p,c];
 *   map_expand: out[p,c] = m[p]*w[c]*(mask[p,c] > 0 if mask);
 *   weighted_channel_sum: sums[(img)*C + c] = sum_p (m ? m[p] : 1)*x[p,c] in fp64 (per image or over all images). */
int vu_scale_by_map(const void* a, int a_ctot, int a_coff, const float* m, void* out, int o_ctot, int o_coff,
                    size_t npix, int C, int accumulate, int dtype, vu_stream_t stream);
int vu_channel_dot(const void* a, int a_ctot, int a_coff, const void* b, int b_ctot, int b_coff, float* out,
                   size_t npix, int C, int dtype, vu_stream_t stream);
int vu_map_expand(const float* m, const float* w, const void* mask, int k_ctot, int k_coff, void* out, int o_ctot,
                  int o_coff, size_t npix, int C, int dtype, vu_stream_t stream);
int vu_weighted_channel_sum(const void* x, int x_ctot, int x_coff, const float* m, int n_img, size_t pix_per_img, int C,
                            int per_image, double* sums, int dtype, vu_stream_t stream);
/* psi branch of the gate: single-channel train-mode BN + sigmoid on an fp32 map (unet_resnet.py:17-21) and its
 * backward (dx w.r.t. the pre-BN map; dgamma[0] / dbeta[0] are accumulated into). stat = {mean, invstd}; sums: fp64[2];
 * gamma / beta are device pointers to the live fp32 parameters (no host round trip inside a step). */
int vu_map_bn_sigmoid_fwd(const float* x, size_t n, const float* gamma, const float* beta, float eps, float momentum,
                          float* running_mean, float* running_var, double* sums, float* stat, float* psi,
                          vu_stream_t stream);
int vu_map_bn_sigmoid_bwd(const float* x, const float* dpsi, const float* psi, const float* stat, const float* gamma, size_t n,
                          double* sums, float* dx, float* dgamma /* += */, float* dbeta /* += */, vu_stream_t stream);
int vu_map_sum(const float* x, size_t n, double* sums, vu_stream_t stream);   /* sums[0] = sum x, sums[1] = sum x^2 */
/* Dense layer on M <= 64 vectors = 1x1 conv (+ train-mode BN over the rows) (+ ReLU) on spatially constant maps or
 * pooled features (z_initial, z_proj, mu/logvar heads). y_raw / mean / invstd are saved for the backward. */
int vu_vec_layer_fwd(const float* z, const float* w, const float* b, int M, int L, int C, const float* gamma,
                     const float* beta, float eps, float momentum, double stat_count, float* running_mean,
                     float* running_var, int relu, float* y_raw, float* y, float* mean, float* invstd, vu_stream_t stream);
int vu_vec_layer_bwd(const float* dy, const float* z, const float* w, const float* y_raw, const float* y, int M, int L,
                     int C, const float* gamma, const float* mean, const float* invstd, int relu, float* dz /* += */,
                     float* dw /* += */, float* db /* += */, float* dgamma /* += */, float* dbeta /* += */,
                     float* scratch /* M*C floats, required when dz != NULL: gradient w.r.t. the pre-BN output */,
                     vu_stream_t stream);
/* z = mu + eps*exp(.5 logvar)*T: dmu += sum_s dz, dlogvar += sum_s dz*eps*.5*exp(.5 logvar)*T. */
int vu_reparam_bwd(const float* dz, const float* eps, const float* logvar, float temperature, float* dmu,
                   float* dlogvar, size_t n_per_sample, int num_samples, vu_stream_t stream);
/* Training path of the ResNet stem on the tensor cores: patch matrix out[n,ho,wo, ci*K*K + r*K + s] (16-bit NHWC with
 * Kp >= Cin*K*K channels, Kp % 8 == 0, zero padded) of an NCHW fp32 image, so that the 7x7 s2 conv is vu_conv2d_fwd
 * (1x1) over it with the filter rows from vu_pack_rows_padded ([O][I] fp32 -> [O][I_pad] 16-bit), and its weight
 * gradient is vu_conv2d_wgrad (1x1) of (patches, dY). */
int vu_im2col_small_cin(const float* x_nchw, void* out, int N, int Cin, int H, int W, int K, int S, int P, int Kp,
                        int dtype, vu_stream_t stream);
int vu_pack_rows_padded(const float* w, void* out, int O, int I, int I_pad, int dtype, vu_stream_t stream);
/* Weight gradient of the ResNet stem conv 7x7 s2 p3 (Cin <= 4): dw [Cout][Cin][7][7] fp32, ACCUMULATED (zero it). */
int vu_stem_wgrad(const float* x_nchw, const void* dy, int dy_ctot, int dy_coff, float* dw, int N, int Cin, int H, int W,
                  int Cout, int dtype, vu_stream_t stream);
/* clip_grad_norm_ + AdamW on flat fp32 buffers (train.py:403-408): out += sum g^2; the step derives the clip
 * coefficient min(1, max_norm / (sqrt(sumsq)*grad_scale + 1e-6)) on the device (grad_sumsq NULL = no clipping). */
int vu_sumsq(const float* g, size_t n, double* out, int zero_first, vu_stream_t stream);
int vu_adamw_step(float* p, const float* g, float* m, float* v, size_t n, float lr, float beta1, float beta2, float eps,
                  float weight_decay, int step, const double* grad_sumsq, float max_norm, float grad_scale,
                  vu_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* VAEUNET_B200_H */
