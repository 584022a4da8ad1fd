// Patch-pipeline kernels (reference utils/data_loading.py:302-446 precompute_all_patches, :580-601 preprocess)
// and the batched form of the tile blending of predict_with_patches (visualize_vae.py:361-384,409):
//   * PIL-exact separable resampling of an 8-bit HWC image with host-built fixed-point coefficient tables (Pillow's
//     ImagingResample: 22-bit coefficients, 8-bit intermediate between the horizontal and the vertical pass) and a
//     nearest-neighbour gather with host-built index tables (ImagingScaleAffine) for the masks;
//   * tile gather: a batch of equal-sized tiles cut out of one device-resident image (uint8 HWC or fp32 CHW) straight into
//     the NCHW fp32 batch the network's first conv reads - no per-tile torch.cat / host staging;
//   * per-tile black-border ratio and lesion flag (is_valid_patch / has_lesion of the reference);
//   * blend_accumulate over a whole tile batch and all MC samples in one launch, output-pixel-centric (each pixel of the
//     batch's bounding box is owned by one thread that walks the tiles in index order: no atomics, the same summation order
//     as one launch per tile), with ONE weight map for all samples.
// All memory-bound streaming kernels; algorithmic bytes = inputs + outputs once.
#include "vu_internal.h"
#include "vu_half.cuh"

namespace vu {

constexpr int kResPrecisionBits = 32 - 8 - 2;      // Pillow's PRECISION_BITS for 8-bit channels

__device__ __forceinline__ uint8_t clip8(int v) {
    v >>= kResPrecisionBits;                       // arithmetic shift, like Pillow's clip8 lookup
    return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

// One pass of Pillow's 8-bit resampling along x (AXIS = 0) or y (AXIS = 1) of an HWC image.
// bounds[o] = {first input index, number of taps}; kk[o * ksize + t] = fixed-point coefficient.
template <int AXIS>
__global__ void __launch_bounds__(256)
resample_u8_kernel(const uint8_t* __restrict__ src, int Hs, int Ws, int C, uint8_t* __restrict__ dst, int Hd, int Wd,
                   const int* __restrict__ bounds, const int* __restrict__ kk, int ksize) {
    const size_t total = (size_t)Hd * Wd * C;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int c = idx % C;
        const int x = (idx / C) % Wd;
        const int y = idx / ((size_t)C * Wd);
        const int o = AXIS == 0 ? x : y;
        const int first = bounds[2 * o], n = bounds[2 * o + 1];
        const int* k = kk + (size_t)o * ksize;
        int ss = 1 << (kResPrecisionBits - 1);
        if (AXIS == 0) {
            const uint8_t* p = src + ((size_t)y * Ws + first) * C + c;
            for (int t = 0; t < n; ++t) ss += (int)p[(size_t)t * C] * k[t];
        } else {
            const uint8_t* p = src + ((size_t)first * Ws + x) * C + c;
            for (int t = 0; t < n; ++t) ss += (int)p[(size_t)t * Ws * C] * k[t];
        }
        dst[idx] = clip8(ss);
    }
}

// dst[y][x][c] = src[yi[y]][xi[x]][c]  (-1 in a table = outside: 0)
__global__ void __launch_bounds__(256)
nearest_u8_kernel(const uint8_t* __restrict__ src, int Hs, int Ws, int C, uint8_t* __restrict__ dst, int Hd, int Wd,
                  const int* __restrict__ yi, const int* __restrict__ xi) {
    const size_t total = (size_t)Hd * Wd * C;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int c = idx % C;
        const int x = (idx / C) % Wd;
        const int y = idx / ((size_t)C * Wd);
        const int sy = yi[y], sx = xi[x];
        dst[idx] = (sy >= 0 && sx >= 0 && sy < Hs && sx < Ws) ? src[((size_t)sy * Ws + sx) * C + c] : (uint8_t)0;
    }
}

// out[t][c][y][x] = scale * src(c, y0[t] + y, x0[t] + x); SRC_U8: src is uint8 HWC, else fp32 CHW. threshold > 0: the output
// is binarised (value > 0 -> 1), the reference's mask conversion (arr > 0).astype(float32).
template <bool SRC_U8>
__global__ void __launch_bounds__(256)
tile_gather_kernel(const void* __restrict__ src_, int C, int H, int W, const int* __restrict__ coords, int T, int ph, int pw,
                   float inv_scale, int binarise, float* __restrict__ out) {
    const size_t per_tile = (size_t)C * ph * pw;
    const size_t total = per_tile * T;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int x = idx % pw;
        const int y = (idx / pw) % ph;
        const int c = (idx / ((size_t)pw * ph)) % C;
        const int t = idx / per_tile;
        const int sy = coords[2 * t] + y, sx = coords[2 * t + 1] + x;
        float v = 0.f;
        if (sy >= 0 && sy < H && sx >= 0 && sx < W) {
            if (SRC_U8) v = (float)static_cast<const uint8_t*>(src_)[((size_t)sy * W + sx) * C + c];
            else v = static_cast<const float*>(src_)[((size_t)c * H + sy) * W + sx];
        }
        if (binarise) v = v > 0.f ? 1.f : 0.f;
        else if (SRC_U8) v = v / inv_scale;        // IEEE division, like numpy's arr.astype(float32) / 255.0
        out[idx] = v;
    }
}

// stats[t] = {fraction of pixels whose channel mean is < black_thr, 1 if any mask pixel > 0.5 else 0}
__global__ void __launch_bounds__(256)
tile_stats_kernel(const float* __restrict__ tiles, int C, int ph, int pw, const float* __restrict__ masks, float black_thr,
                  float* __restrict__ stats) {
    const int t = blockIdx.x;
    const size_t npix = (size_t)ph * pw;
    const float* tp = tiles + (size_t)t * C * npix;
    unsigned int black = 0, lesion = 0;
    for (size_t i = threadIdx.x; i < npix; i += blockDim.x) {
        float s = 0.f;
        for (int c = 0; c < C; ++c) s += tp[(size_t)c * npix + i];      // torch: patch.mean(dim=0) sums the channels in order
        if (s / (float)C < black_thr) ++black;
        if (masks != nullptr && masks[(size_t)t * npix + i] > 0.5f) lesion = 1;
    }
    black = __reduce_add_sync(0xffffffffu, black);
    lesion = __reduce_or_sync(0xffffffffu, lesion);
    __shared__ unsigned int sb[8], sl[8];
    if ((threadIdx.x & 31) == 0) { sb[threadIdx.x >> 5] = black; sl[threadIdx.x >> 5] = lesion; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int b = 0, l = 0;
        for (int w = 0; w < 8; ++w) { b += sb[w]; l |= sl[w]; }
        stats[2 * t] = (float)b / (float)npix;
        stats[2 * t + 1] = l ? 1.f : 0.f;
    }
}

__device__ __forceinline__ float ramp01(int i, int n) {       // torch.linspace(0, 1, n)[i] in fp32 (steps = n)
    if (n <= 1) return 0.f;
    const float step = 1.0f / (float)(n - 1);
    // torch.linspace fills the first half forward and the second half backward from the end point
    return (i < n / 2) ? (float)i * step : 1.0f - (float)(n - 1 - i) * step;
}

// One launch for a batch of T tiles x Ns samples. info[t] = {sh, sw, ti, tj}. Thread = one pixel of the batch's bounding box
// [y0, y1) x [x0, x1): for every tile (in index order) that covers it, acc[n] += tile[t][n] * w and wsum += w.
__global__ void __launch_bounds__(256)
blend_accumulate_batch_kernel(const float* __restrict__ tiles, const int* __restrict__ info, int T, int Ns, int ph, int pw,
                              int nth, int ntw, int ov, float* __restrict__ acc, float* __restrict__ wsum, int H, int W,
                              int y0, int y1, int x0, int x1) {
    const int bw = x1 - x0;
    const size_t total = (size_t)(y1 - y0) * bw;
    const size_t plane = (size_t)H * W, tplane = (size_t)ph * pw;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int gx = x0 + (int)(idx % bw), gy = y0 + (int)(idx / bw);
        const size_t o = (size_t)gy * W + gx;
        float ws = wsum[o];                        // running sum in tile order: the same fp32 order as one launch per tile
        bool any = false;
        for (int t = 0; t < T; ++t) {
            const int sh = info[4 * t], sw = info[4 * t + 1], ti = info[4 * t + 2], tj = info[4 * t + 3];
            const int y = gy - sh, x = gx - sw;
            if (y < 0 || y >= ph || x < 0 || x >= pw) continue;
            float w = 1.0f;
            if (ph > 2 * ov) {                     // visualize_vae.py:364-378 (axis 2 first, then axis 3)
                if (ti > 0 && y < ov) w *= ramp01(y, ov);
                if (ti < nth - 1 && y >= ph - ov) w *= (1.0f - ramp01(y - (ph - ov), ov));
            }
            if (pw > 2 * ov) {
                if (tj > 0 && x < ov) w *= ramp01(x, ov);
                if (tj < ntw - 1 && x >= pw - ov) w *= (1.0f - ramp01(x - (pw - ov), ov));
            }
            const float* tp = tiles + ((size_t)t * Ns) * tplane + (size_t)y * pw + x;
            for (int n = 0; n < Ns; ++n) acc[(size_t)n * plane + o] += tp[(size_t)n * tplane] * w;
            ws += w;
            any = true;
        }
        if (any) wsum[o] = ws;
    }
}

// out[n][i] = acc[n][i] / (wsum[i] + 1e-8): one weight map for all samples
__global__ void __launch_bounds__(256)
blend_finalize_shared_kernel(const float* __restrict__ acc, const float* __restrict__ wsum, float* __restrict__ out, size_t plane,
                             int Ns) {
    const size_t total = plane * Ns;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x)
        out[idx] = acc[idx] / (wsum[idx % plane] + 1e-8f);
}

static int tl_grid(size_t n, int per_block = 256 * 4) {
    size_t g = (n + per_block - 1) / per_block;
    const size_t cap = (size_t)num_sms() * 16;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

}  // namespace vu

using namespace vu;

extern "C" int vu_resample_u8(const uint8_t* src, int Hs, int Ws, int C, uint8_t* dst, int Hd, int Wd, int axis,
                              const int* bounds, const int* kk, int ksize, vu_stream_t st) {
    if (!src || !dst || !bounds || !kk || C < 1 || ksize < 1) return set_error(VU_ERR_INVALID, "resample_u8: bad args");
    if ((axis == 0 && Hd != Hs) || (axis == 1 && Wd != Ws) || (axis != 0 && axis != 1))
        return set_error(VU_ERR_INVALID, "resample_u8: a pass changes one axis (axis 0 = x: Hd == Hs; axis 1 = y: Wd == Ws)");
    const size_t total = (size_t)Hd * Wd * C;
    if (axis == 0)
        resample_u8_kernel<0><<<tl_grid(total), 256, 0, (cudaStream_t)st>>>(src, Hs, Ws, C, dst, Hd, Wd, bounds, kk, ksize);
    else
        resample_u8_kernel<1><<<tl_grid(total), 256, 0, (cudaStream_t)st>>>(src, Hs, Ws, C, dst, Hd, Wd, bounds, kk, ksize);
    VU_CHECK_LAUNCH("resample_u8");
    return 0;
}

extern "C" int vu_nearest_u8(const uint8_t* src, int Hs, int Ws, int C, uint8_t* dst, int Hd, int Wd, const int* yi,
                             const int* xi, vu_stream_t st) {
    if (!src || !dst || !yi || !xi || C < 1) return set_error(VU_ERR_INVALID, "nearest_u8: bad args");
    nearest_u8_kernel<<<tl_grid((size_t)Hd * Wd * C), 256, 0, (cudaStream_t)st>>>(src, Hs, Ws, C, dst, Hd, Wd, yi, xi);
    VU_CHECK_LAUNCH("nearest_u8");
    return 0;
}

extern "C" int vu_tile_gather(const void* src, int src_is_u8, int C, int H, int W, const int* coords, int T, int ph, int pw,
                              float divisor, int binarise, float* out, vu_stream_t st) {
    if (!src || !coords || !out || C < 1 || T < 1 || ph < 1 || pw < 1) return set_error(VU_ERR_INVALID, "tile_gather: bad args");
    if (divisor == 0.f) return set_error(VU_ERR_INVALID, "tile_gather: divisor 0");
    const size_t total = (size_t)T * C * ph * pw;
    if (src_is_u8)
        tile_gather_kernel<true><<<tl_grid(total), 256, 0, (cudaStream_t)st>>>(src, C, H, W, coords, T, ph, pw, divisor, binarise, out);
    else
        tile_gather_kernel<false><<<tl_grid(total), 256, 0, (cudaStream_t)st>>>(src, C, H, W, coords, T, ph, pw, divisor, binarise, out);
    VU_CHECK_LAUNCH("tile_gather");
    return 0;
}

extern "C" int vu_tile_stats(const float* tiles, int T, int C, int ph, int pw, const float* masks, float black_thr, float* stats,
                             vu_stream_t st) {
    if (!tiles || !stats || T < 1) return set_error(VU_ERR_INVALID, "tile_stats: bad args");
    tile_stats_kernel<<<T, 256, 0, (cudaStream_t)st>>>(tiles, C, ph, pw, masks, black_thr, stats);
    VU_CHECK_LAUNCH("tile_stats");
    return 0;
}

extern "C" int vu_blend_accumulate_batch(const float* tiles, const int* info, int T, int Ns, int ph, int pw, int nth, int ntw,
                                         int overlap, float* acc, float* wsum, int H, int W, int y0, int y1, int x0, int x1,
                                         vu_stream_t st) {
    if (!tiles || !info || !acc || !wsum || T < 1 || Ns < 1) return set_error(VU_ERR_INVALID, "blend_accumulate_batch: bad args");
    if (y0 < 0 || x0 < 0 || y1 > H || x1 > W || y1 <= y0 || x1 <= x0) return set_error(VU_ERR_INVALID, "blend_accumulate_batch: bounding box outside the image");
    if (overlap < 0) return set_error(VU_ERR_INVALID, "blend_accumulate_batch: negative overlap");
    const size_t total = (size_t)(y1 - y0) * (x1 - x0);
    blend_accumulate_batch_kernel<<<tl_grid(total, 256), 256, 0, (cudaStream_t)st>>>(tiles, info, T, Ns, ph, pw, nth, ntw, overlap, acc,
                                                                                     wsum, H, W, y0, y1, x0, x1);
    VU_CHECK_LAUNCH("blend_accumulate_batch");
    return 0;
}

extern "C" int vu_blend_finalize_shared(const float* acc, const float* wsum, float* out, size_t plane, int Ns, vu_stream_t st) {
    if (!acc || !wsum || !out || Ns < 1) return set_error(VU_ERR_INVALID, "blend_finalize_shared: bad args");
    blend_finalize_shared_kernel<<<tl_grid(plane * Ns), 256, 0, (cudaStream_t)st>>>(acc, wsum, out, plane, Ns);
    VU_CHECK_LAUNCH("blend_finalize_shared");
    return 0;
}
