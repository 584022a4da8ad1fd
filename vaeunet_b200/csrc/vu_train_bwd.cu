// Backward passes of the memory-bound forward ops (training step): max-pool, bilinear
// align_corners resize (16-bit NHWC maps and fp32 single-channel maps), and the per-pixel pieces of the
// attention gate / output conv (scale by a map, channel dot product, outer product with a vector,
// map-weighted channel sums, per-image channel sums). Gather formulations: deterministic, no 16-bit atomics.
#include <cstdlib>
#include "vu_half.cuh"
#include "vu_internal.h"

namespace vu {

static int tb_grid(size_t items, int per_block = 256, int mult = 8) {
    size_t g = (items + per_block - 1) / per_block;
    const size_t cap = (size_t)num_sms() * mult;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

// ---------------------------------------------------------------- max-pool backward
// dx[n,ih,iw,c] = sum over the windows (ho,wo) containing (ih,iw) whose FIRST maximum (row-major scan, as ATen)
// is at (ih,iw) of dy[n,ho,wo,c].
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const uint16_t* __restrict__ x, int x_ctot, int x_coff, const uint16_t* __restrict__ dy, int dy_ctot,
                   int dy_coff, uint16_t* __restrict__ dx, int dx_ctot, int dx_coff, int N, int H, int W, int C8, int k,
                   int s, int p, int Ho, int Wo, int dtype) {
    const size_t total = (size_t)N * H * W * C8;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int c8 = idx % C8;
        const int iw = (idx / C8) % W;
        const int ih = (idx / ((size_t)C8 * W)) % H;
        const int n = idx / ((size_t)C8 * W * H);
        const uint16_t* xb = x + ((size_t)n * H * W) * x_ctot + x_coff;
        float me[8], acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        unpack8(__ldg(reinterpret_cast<const uint4*>(xb + ((size_t)ih * W + iw) * x_ctot) + c8), me, dtype);
        const int ho_lo = max(0, (ih + p - k + s) / s), ho_hi = min(Ho - 1, (ih + p) / s);
        const int wo_lo = max(0, (iw + p - k + s) / s), wo_hi = min(Wo - 1, (iw + p) / s);
        for (int ho = ho_lo; ho <= ho_hi; ++ho)
            for (int wo = wo_lo; wo <= wo_hi; ++wo) {
                bool win[8] = {true, true, true, true, true, true, true, true};
                for (int r = 0; r < k; ++r) {
                    const int h2 = ho * s - p + r;
                    if (h2 < 0 || h2 >= H) continue;
                    for (int q = 0; q < k; ++q) {
                        const int w2 = wo * s - p + q;
                        if (w2 < 0 || w2 >= W || (h2 == ih && w2 == iw)) continue;
                        float o[8];
                        unpack8(__ldg(reinterpret_cast<const uint4*>(xb + ((size_t)h2 * W + w2) * x_ctot) + c8), o, dtype);
                        const bool before = (h2 < ih) || (h2 == ih && w2 < iw);
#pragma unroll
                        for (int j = 0; j < 8; ++j) win[j] = win[j] && (before ? me[j] > o[j] : me[j] >= o[j]);
                    }
                }
                float g[8];
                unpack8(__ldg(reinterpret_cast<const uint4*>(dy + (((size_t)n * Ho + ho) * Wo + wo) * dy_ctot + dy_coff) + c8), g, dtype);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] += win[j] ? g[j] : 0.f;
            }
        *(reinterpret_cast<uint4*>(dx + (((size_t)n * H + ih) * W + iw) * dx_ctot + dx_coff) + c8) = pack8(acc, dtype);
    }
}

// ---------------------------------------------------------------- max-pool with saved argmax (training fast path)
// Forward also writes, per output element, the position r*k+q of the FIRST maximum of its window (row-major scan, as
// ATen); backward is then a pure gather: input pixel (ih,iw) sums dy over the <= 4 windows whose code points at it.
__global__ void __launch_bounds__(256)
maxpool_idx_fwd_kernel(const uint16_t* __restrict__ x, int x_ctot, int x_coff, uint16_t* __restrict__ y, int y_ctot,
                       int y_coff, uint8_t* __restrict__ code, int N, int H, int W, int C8, int k, int s, int p, int Ho,
                       int Wo, int dtype) {
    const size_t total = (size_t)N * Ho * Wo * C8;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int c8 = idx % C8;
        const int wo = (idx / C8) % Wo;
        const int ho = (idx / ((size_t)C8 * Wo)) % Ho;
        const int n = idx / ((size_t)C8 * Wo * Ho);
        float m[8];
        uint32_t cd[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { m[j] = -INFINITY; cd[j] = 0xFFu; }
        for (int r = 0; r < k; ++r) {
            const int ih = ho * s - p + r;
            if (ih < 0 || ih >= H) continue;
            for (int q = 0; q < k; ++q) {
                const int iw = wo * s - p + q;
                if (iw < 0 || iw >= W) continue;
                float f[8];
                unpack8(__ldg(reinterpret_cast<const uint4*>(x + (((size_t)n * H + ih) * W + iw) * x_ctot + x_coff) + c8), f, dtype);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (f[j] > m[j] || cd[j] == 0xFFu) { m[j] = f[j]; cd[j] = (uint32_t)(r * k + q); }
            }
        }
        const size_t o = ((size_t)n * Ho + ho) * Wo + wo;
        *(reinterpret_cast<uint4*>(y + o * y_ctot + y_coff) + c8) = pack8(m, dtype);
        uint2 pk;
        pk.x = cd[0] | (cd[1] << 8) | (cd[2] << 16) | (cd[3] << 24);
        pk.y = cd[4] | (cd[5] << 8) | (cd[6] << 16) | (cd[7] << 24);
        *(reinterpret_cast<uint2*>(code + o * (size_t)(C8 * 8)) + c8) = pk;
    }
}

template <typename I>   // 32-bit index arithmetic whenever the element count allows
__global__ void __launch_bounds__(256)
maxpool_idx_bwd_kernel(const uint8_t* __restrict__ code, const uint16_t* __restrict__ dy, int dy_ctot, int dy_coff,
                       uint16_t* __restrict__ dx, int dx_ctot, int dx_coff, int N, int H, int W, int C8, int k, int s, int p,
                       int Ho, int Wo, int dtype) {
    const I total = (I)N * H * W * C8;
    for (I idx = blockIdx.x * (I)blockDim.x + threadIdx.x; idx < total; idx += (I)gridDim.x * blockDim.x) {
        const int c8 = idx % C8;
        const int iw = (idx / C8) % W;
        const int ih = (idx / ((I)C8 * W)) % H;
        const int n = idx / ((I)C8 * W * H);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const int ho_lo = max(0, (ih + p - k + s) / s), ho_hi = min(Ho - 1, (ih + p) / s);
        const int wo_lo = max(0, (iw + p - k + s) / s), wo_hi = min(Wo - 1, (iw + p) / s);
        for (int ho = ho_lo; ho <= ho_hi; ++ho)
            for (int wo = wo_lo; wo <= wo_hi; ++wo) {
                const uint32_t me = (uint32_t)((ih - (ho * s - p)) * k + (iw - (wo * s - p)));
                const size_t o = ((size_t)n * Ho + ho) * Wo + wo;
                const uint2 pk = __ldg(reinterpret_cast<const uint2*>(code + o * (size_t)(C8 * 8)) + c8);
                float g[8];
                unpack8(__ldg(reinterpret_cast<const uint4*>(dy + o * dy_ctot + dy_coff) + c8), g, dtype);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t cj = ((j < 4 ? pk.x : pk.y) >> (8 * (j & 3))) & 0xFFu;
                    acc[j] += (cj == me) ? g[j] : 0.f;
                }
            }
        *(reinterpret_cast<uint4*>(dx + (((size_t)n * H + ih) * W + iw) * dx_ctot + dx_coff) + c8) = pack8(acc, dtype);
    }
}

// ResNet's pool (k 3, s 2, p 1) on a power-of-two channel count: one block row per input row (no div/mod), the eight window
// codes of a chunk compared in two SIMD byte compares, the matching dy halves selected by AND masks before the fp32 sum
// (same sums as the generic kernel: the non-matching terms are exact zeros either way).
template <int C8>
__global__ void __launch_bounds__(256)
maxpool_idx_bwd_k3s2_kernel(const uint8_t* __restrict__ code, const uint16_t* __restrict__ dy, int dy_ctot, int dy_coff,
                            uint16_t* __restrict__ dx, int dx_ctot, int dx_coff, int H, int W, int Ho, int Wo, int dtype) {
    const int ih = blockIdx.y, n = blockIdx.z;
    const int t = blockIdx.x * 256 + threadIdx.x;
    const int iw = t / C8, c8 = t % C8;
    if (iw >= W) return;
    const int ho_lo = ih >> 1, ho_hi = min(Ho - 1, (ih + 1) >> 1);
    const int wo_lo = iw >> 1, wo_hi = min(Wo - 1, (iw + 1) >> 1);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int a = 0; a < 2; ++a) {
        const int ho = ho_lo + a;
        if (ho > ho_hi) break;
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const int wo = wo_lo + b;
            if (wo > wo_hi) break;
            const uint32_t me = (uint32_t)((ih - (ho * 2 - 1)) * 3 + (iw - (wo * 2 - 1)));
            const uint32_t me4 = me * 0x01010101u;
            const size_t o = ((size_t)n * Ho + ho) * Wo + wo;
            const uint2 pk = __ldg(reinterpret_cast<const uint2*>(code + o * (size_t)(C8 * 8)) + c8);
            uint4 v = __ldg(reinterpret_cast<const uint4*>(dy + o * dy_ctot + dy_coff) + c8);
            const uint32_t m0 = __vcmpeq4(pk.x, me4), m1 = __vcmpeq4(pk.y, me4);
            v.x &= __byte_perm(m0, 0, 0x1100); v.y &= __byte_perm(m0, 0, 0x3322);
            v.z &= __byte_perm(m1, 0, 0x1100); v.w &= __byte_perm(m1, 0, 0x3322);
            float g[8];
            unpack8(v, g, dtype);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += g[j];
        }
    }
    *(reinterpret_cast<uint4*>(dx + (((size_t)n * H + ih) * W + iw) * dx_ctot + dx_coff) + c8) = pack8(acc, dtype);
}

// ---------------------------------------------------------------- bilinear (align_corners=True) backward
// Adjoint of vu_bilinear_ac_nhwc as a gather: source pixel (hs,ws) collects w(ho;hs) * w(wo;ws) * dy[ho,wo] over the
// few destination rows / columns whose interpolation footprint contains it (same fp32 index arithmetic as forward).
__device__ __forceinline__ float lerp_weight(int d, int s, float scale, int in_size) {
    const float r = scale * d;
    const int i0 = (int)r;
    const int ip = (i0 < in_size - 1) ? 1 : 0;
    const float l1 = r - i0;
    float w = 0.f;
    if (i0 == s) w += 1.0f - l1;
    if (i0 + ip == s) w += l1;
    return w;
}
__device__ __forceinline__ void dst_range(int s, float scale, int out_size, int& lo, int& hi) {
    if (scale <= 0.f) { lo = 0; hi = out_size - 1; return; }
    lo = max(0, (int)floorf((s - 1) / scale) - 1);
    hi = min(out_size - 1, (int)ceilf((s + 1) / scale) + 1);
}

__global__ void __launch_bounds__(256)
bilinear_bwd_kernel(const uint16_t* __restrict__ dy, int dy_ctot, int dy_coff, uint16_t* __restrict__ dx, int dx_ctot,
                    int dx_coff, int N, int Hs, int Ws, int Ho, int Wo, int C8, float sh, float sw, int accumulate,
                    int dtype) {
    const size_t total = (size_t)N * Hs * Ws * C8;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int c8 = idx % C8;
        const int ws = (idx / C8) % Ws;
        const int hs = (idx / ((size_t)C8 * Ws)) % Hs;
        const int n = idx / ((size_t)C8 * Ws * Hs);
        int hlo, hhi, wlo, whi;
        dst_range(hs, sh, Ho, hlo, hhi);
        dst_range(ws, sw, Wo, wlo, whi);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int ho = hlo; ho <= hhi; ++ho) {
            const float wh = lerp_weight(ho, hs, sh, Hs);
            if (wh == 0.f) continue;
            for (int wo = wlo; wo <= whi; ++wo) {
                const float ww = lerp_weight(wo, ws, sw, Ws);
                if (ww == 0.f) continue;
                float g[8];
                unpack8(__ldg(reinterpret_cast<const uint4*>(dy + (((size_t)n * Ho + ho) * Wo + wo) * dy_ctot + dy_coff) + c8), g, dtype);
                const float w = wh * ww;
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fmaf(w, g[j], acc[j]);
            }
        }
        uint4* dp = reinterpret_cast<uint4*>(dx + (((size_t)n * Hs + hs) * Ws + ws) * dx_ctot + dx_coff) + c8;
        if (accumulate) {
            float o[8];
            unpack8(*dp, o, dtype);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += o[j];
        }
        *dp = pack8(acc, dtype);
    }
}

// Fast path for the x2 upsamples of the model (scale >= 0.4). The destination rows / columns whose footprint contains a
// source pixel are CONSECUTIVE and at most kBilTaps = 5 (d*scale in (s-1, s+1): an interval of length 2/scale <= 5), so the
// taps are d0 .. d0+4 with d0 the first non-zero weight (same fp32 index arithmetic as the forward kernel); zero rows are
// skipped, the five column loads of a row are issued together.
constexpr int kBilTaps = 5;
__device__ __forceinline__ int lerp_taps(int s, float scale, int in_size, int out_size, float (&w)[kBilTaps]) {
    int lo, hi;
    dst_range(s, scale, out_size, lo, hi);
    int d0 = lo;
    while (d0 < hi && lerp_weight(d0, s, scale, in_size) == 0.f) ++d0;
#pragma unroll
    for (int k = 0; k < kBilTaps; ++k) w[k] = (d0 + k <= hi) ? lerp_weight(d0 + k, s, scale, in_size) : 0.f;
    return d0;
}

template <typename I>
__global__ void __launch_bounds__(256)
bilinear_bwd_taps_kernel(const uint16_t* __restrict__ dy, int dy_ctot, int dy_coff, uint16_t* __restrict__ dx, int dx_ctot,
                         int dx_coff, int N, int Hs, int Ws, int Ho, int Wo, int C8, float sh, float sw, int accumulate,
                         int dtype) {
    const I total = (I)N * Hs * Ws * C8;
    for (I idx = blockIdx.x * (I)blockDim.x + threadIdx.x; idx < total; idx += (I)gridDim.x * blockDim.x) {
        const int c8 = idx % C8;
        const int ws = (idx / C8) % Ws;
        const int hs = (idx / ((I)C8 * Ws)) % Hs;
        const int n = idx / ((I)C8 * Ws * Hs);
        float rw[kBilTaps], cw[kBilTaps];
        const int r0 = lerp_taps(hs, sh, Hs, Ho, rw);
        const int c0 = lerp_taps(ws, sw, Ws, Wo, cw);
        const uint16_t* base = dy + ((size_t)n * Ho * Wo + c0) * dy_ctot + dy_coff + c8 * 8;
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int i = 0; i < kBilTaps; ++i) {
            if (rw[i] == 0.f) continue;
            const uint16_t* row = base + (size_t)(r0 + i) * Wo * dy_ctot;
            uint4 v[kBilTaps];
#pragma unroll
            for (int k = 0; k < kBilTaps; ++k)
                v[k] = (cw[k] != 0.f) ? __ldg(reinterpret_cast<const uint4*>(row + (size_t)k * dy_ctot)) : make_uint4(0, 0, 0, 0);
            float t[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
            for (int k = 0; k < kBilTaps; ++k) {
                float g[8];
                unpack8(v[k], g, dtype);
#pragma unroll
                for (int q = 0; q < 8; ++q) t[q] = fmaf(cw[k], g[q], t[q]);
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) acc[q] = fmaf(rw[i], t[q], acc[q]);
        }
        uint4* dp = reinterpret_cast<uint4*>(dx + (((size_t)n * Hs + hs) * Ws + ws) * dx_ctot + dx_coff) + c8;
        if (accumulate) {
            float o[8];
            unpack8(*dp, o, dtype);
#pragma unroll
            for (int q = 0; q < 8; ++q) acc[q] += o[q];
        }
        *dp = pack8(acc, dtype);
    }
}

// The same gather with the tap tables built ONCE per run of source pixels: a warp owns RUN = min(32, 256 / C8) consecutive
// source pixels of one source row, lane p computes the first destination column and the five column weights of pixel p into
// shared memory (lerp_taps costs ~300 instructions: per pixel now, not per 16-byte chunk), the row taps are warp-uniform.
// Lanes then walk the flattened (pixel, chunk) list with compile-time shifts. Arithmetic and summation order are those of
// bilinear_bwd_taps_kernel: results are bit-identical.
template <int C8>
__global__ void __launch_bounds__(128)
bilinear_bwd_run_kernel(const uint16_t* __restrict__ dy, int dy_ctot, int dy_coff, uint16_t* __restrict__ dx, int dx_ctot,
                        int dx_coff, int Hs, int Ws, int Ho, int Wo, float sh, float sw, int accumulate, int dtype,
                        int runs_per_row, int num_runs) {
    constexpr int RUN = (256 / C8 < 32) ? 256 / C8 : 32;
    constexpr int PASSES = RUN * C8 / 32;
    __shared__ float s_cw[4][32][kBilTaps];
    __shared__ int s_c0[4][32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int run = blockIdx.x * 4 + warp;
    if (run >= num_runs) return;                 // warp-uniform; only __syncwarp below
    const int n = blockIdx.y;
    const int hs = run / runs_per_row;
    const int ws0 = (run - hs * runs_per_row) * RUN;
    const int npx = min(RUN, Ws - ws0);
    float rw[kBilTaps];
    const int r0 = lerp_taps(hs, sh, Hs, Ho, rw);
    if (lane < RUN) {
        float cw[kBilTaps];
        s_c0[warp][lane] = lerp_taps(ws0 + min(lane, npx - 1), sw, Ws, Wo, cw);
#pragma unroll
        for (int k = 0; k < kBilTaps; ++k) s_cw[warp][lane][k] = cw[k];
    }
    __syncwarp();
    const uint16_t* img = dy + (size_t)n * Ho * Wo * dy_ctot + dy_coff;
    uint16_t* orow = dx + (((size_t)n * Hs + hs) * Ws + ws0) * dx_ctot + dx_coff;
#pragma unroll 1
    for (int pass = 0; pass < PASSES; ++pass) {
        const int idx = pass * 32 + lane, px = idx / C8, c8 = idx % C8;
        float cw[kBilTaps];
#pragma unroll
        for (int k = 0; k < kBilTaps; ++k) cw[k] = s_cw[warp][px][k];
        const uint16_t* base = img + (size_t)s_c0[warp][px] * dy_ctot + c8 * 8;
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int i = 0; i < kBilTaps; ++i) {
            if (rw[i] == 0.f) continue;            // warp-uniform
            const uint16_t* row = base + (size_t)(r0 + i) * Wo * dy_ctot;
            uint4 v[kBilTaps];
#pragma unroll
            for (int k = 0; k < kBilTaps; ++k)
                v[k] = (cw[k] != 0.f) ? __ldg(reinterpret_cast<const uint4*>(row + (size_t)k * dy_ctot)) : make_uint4(0, 0, 0, 0);
            float t[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
            for (int k = 0; k < kBilTaps; ++k) {
                float g[8];
                unpack8(v[k], g, dtype);
#pragma unroll
                for (int q = 0; q < 8; ++q) t[q] = fmaf(cw[k], g[q], t[q]);
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) acc[q] = fmaf(rw[i], t[q], acc[q]);
        }
        if (px < npx) {
            uint4* dp = reinterpret_cast<uint4*>(orow + (size_t)px * dx_ctot) + c8;
            if (accumulate) {
                float o[8];
                unpack8(*dp, o, dtype);
#pragma unroll
                for (int q = 0; q < 8; ++q) acc[q] += o[q];
            }
            *dp = pack8(acc, dtype);
        }
    }
}

template <int C8>
static void launch_bilinear_bwd_run(const uint16_t* dy, int dy_ctot, int dy_coff, uint16_t* dx, int dx_ctot, int dx_coff, int N,
                                    int Hs, int Ws, int Ho, int Wo, float sh, float sw, int accumulate, int dtype,
                                    cudaStream_t st) {
    constexpr int RUN = (256 / C8 < 32) ? 256 / C8 : 32;
    const int runs_per_row = (Ws + RUN - 1) / RUN, num_runs = Hs * runs_per_row;
    bilinear_bwd_run_kernel<C8><<<dim3((num_runs + 3) / 4, N), 128, 0, st>>>(dy, dy_ctot, dy_coff, dx, dx_ctot, dx_coff, Hs, Ws, Ho,
                                                                           Wo, sh, sw, accumulate, dtype, runs_per_row, num_runs);
}

__global__ void __launch_bounds__(256)
resize_map_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, int N, int Hs, int Ws, int Ho, int Wo,
                      float sh, float sw) {
    const size_t total = (size_t)N * Hs * Ws;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int ws = idx % Ws, hs = (idx / Ws) % Hs, n = idx / ((size_t)Ws * Hs);
        int hlo, hhi, wlo, whi;
        dst_range(hs, sh, Ho, hlo, hhi);
        dst_range(ws, sw, Wo, wlo, whi);
        float acc = 0.f;
        for (int ho = hlo; ho <= hhi; ++ho) {
            const float wh = lerp_weight(ho, hs, sh, Hs);
            if (wh == 0.f) continue;
            for (int wo = wlo; wo <= whi; ++wo)
                acc = fmaf(wh * lerp_weight(wo, ws, sw, Ws), __ldg(dy + ((size_t)n * Ho + ho) * Wo + wo), acc);
        }
        dx[idx] = acc;
    }
}

// ---------------------------------------------------------------- per-pixel pieces (attention gate, output conv)
// out[p,c] = a[p / rep ...] * m[p]: a has Na = N / rep images (MC / none in training: rep = 1), m fp32 [N*HW]
__global__ void __launch_bounds__(256)
scale_by_map_kernel(const uint16_t* __restrict__ a, int a_ctot, int a_coff, const float* __restrict__ m,
                    uint16_t* __restrict__ out, int o_ctot, int o_coff, size_t npix, int C8, int accumulate, int dtype) {
    const size_t total = npix * C8;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int c8 = idx % C8;
        const size_t p = idx / C8;
        float f[8];
        unpack8(__ldg(reinterpret_cast<const uint4*>(a + p * a_ctot + a_coff) + c8), f, dtype);
        const float s = __ldg(m + p);
        uint4* op = reinterpret_cast<uint4*>(out + p * o_ctot + o_coff) + c8;
        float o[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (accumulate) unpack8(*op, o, dtype);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = fmaf(f[j], s, o[j]);
        *op = pack8(o, dtype);
    }
}

// out[p] = sum_c a[p,c] * b[p,c]   (8 lanes per pixel)
__global__ void __launch_bounds__(256)
channel_dot_kernel(const uint16_t* __restrict__ a, int a_ctot, int a_coff, const uint16_t* __restrict__ b, int b_ctot,
                   int b_coff, float* __restrict__ out, size_t npix, int C8, int dtype) {
    const int sub = threadIdx.x & 7;
    const size_t gid = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 3;
    const size_t gstride = ((size_t)gridDim.x * blockDim.x) >> 3;
    const size_t gl = (threadIdx.x & 31) >> 3;
    for (size_t p_ = gid; p_ - gl < npix; p_ += gstride) {
        const bool valid = p_ < npix;
        const size_t p = valid ? p_ : npix - 1;
        float d = 0.f;
        for (int c8 = sub; c8 < C8; c8 += 8) {
            float fa[8], fb[8];
            unpack8(__ldg(reinterpret_cast<const uint4*>(a + p * a_ctot + a_coff) + c8), fa, dtype);
            unpack8(__ldg(reinterpret_cast<const uint4*>(b + p * b_ctot + b_coff) + c8), fb, dtype);
#pragma unroll
            for (int j = 0; j < 8; ++j) d = fmaf(fa[j], fb[j], d);
        }
        d += __shfl_xor_sync(0xffffffffu, d, 4, 8);
        d += __shfl_xor_sync(0xffffffffu, d, 2, 8);
        d += __shfl_xor_sync(0xffffffffu, d, 1, 8);
        if (valid && sub == 0) out[p] = d;
    }
}

// out[p,c] = m[p] * w[c] * (mask[p,c] > 0 if mask)      (d of a C->1 1x1 conv w.r.t. its input, through a ReLU)
__global__ void __launch_bounds__(256)
map_expand_kernel(const float* __restrict__ m, const float* __restrict__ w, const uint16_t* __restrict__ mask, int k_ctot,
                  int k_coff, uint16_t* __restrict__ out, int o_ctot, int o_coff, size_t npix, int C8, int dtype) {
    const size_t total = npix * C8;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int c8 = idx % C8;
        const size_t p = idx / C8;
        const float s = __ldg(m + p);
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = s * __ldg(w + c8 * 8 + j);
        if (mask != nullptr) {
            float k[8];
            unpack8(__ldg(reinterpret_cast<const uint4*>(mask + p * k_ctot + k_coff) + c8), k, dtype);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = k[j] > 0.f ? o[j] : 0.f;
        }
        *(reinterpret_cast<uint4*>(out + p * o_ctot + o_coff) + c8) = pack8(o, dtype);
    }
}

// sums[img*C + c] += sum over the pixels p of image img of (m ? m[p] : 1) * x[p,c]; per_image = 0 folds all images
// (d of the C->1 conv weight; global-average-pool; gradient of a spatially constant map).
__global__ void __launch_bounds__(256)
weighted_channel_sum_kernel(const uint16_t* __restrict__ x, int x_ctot, int x_coff, const float* __restrict__ m,
                            size_t pix_per_img, int C8, int per_image, int blocks_per_img, double* __restrict__ sums,
                            int dtype) {
    extern __shared__ float sm[];
    const int lanes = 256 / C8;
    const int c8 = threadIdx.x % C8, pl = threadIdx.x / C8;
    const int C = C8 * 8;
    const int img = blockIdx.x / blocks_per_img, bi = blockIdx.x % blocks_per_img;
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (pl < lanes) {
        for (size_t q = (size_t)bi * lanes + pl; q < pix_per_img; q += (size_t)blocks_per_img * lanes) {
            const size_t p = (size_t)img * pix_per_img + q;
            float f[8];
            unpack8(__ldg(reinterpret_cast<const uint4*>(x + p * x_ctot + x_coff) + c8), f, dtype);
            const float s = m ? __ldg(m + p) : 1.0f;
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] = fmaf(s, f[j], a[j]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) sm[pl * C + c8 * 8 + j] = a[j];
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += 256) {
        float t = 0.f;
        for (int l = 0; l < lanes; ++l) t += sm[l * C + c];
        atomicAdd(&sums[(per_image ? (size_t)img * C : 0) + c], (double)t);
    }
}

}  // namespace vu

using namespace vu;

#define VU_TB_ALIGN(cond, what) if (cond) return set_error(VU_ERR_ALIGN, what ": channel counts/offsets must be multiples of 8")

extern "C" int vu_maxpool_bwd(const void* x, const void* dy, void* dx, int N, int H, int W, int C, int x_ctot, int x_coff,
                              int dy_ctot, int dy_coff, int dx_ctot, int dx_coff, int k, int s, int p, int dtype,
                              vu_stream_t st) {
    if (!x || !dy || !dx) return set_error(VU_ERR_INVALID, "maxpool_bwd: null");
    VU_TB_ALIGN((C | x_ctot | x_coff | dy_ctot | dy_coff | dx_ctot | dx_coff) % 8, "maxpool_bwd");
    const int Ho = (H + 2 * p - k) / s + 1, Wo = (W + 2 * p - k) / s + 1;
    maxpool_bwd_kernel<<<tb_grid((size_t)N * H * W * (C / 8)), 256, 0, (cudaStream_t)st>>>(
        (const uint16_t*)x, x_ctot, x_coff, (const uint16_t*)dy, dy_ctot, dy_coff, (uint16_t*)dx, dx_ctot, dx_coff, N, H, W,
        C / 8, k, s, p, Ho, Wo, dtype);
    VU_CHECK_LAUNCH("maxpool_bwd");
    return 0;
}

extern "C" int vu_maxpool_idx_fwd(const void* x, void* y, void* code, int N, int H, int W, int C, int x_ctot, int x_coff,
                                  int y_ctot, int y_coff, int k, int s, int p, int dtype, vu_stream_t st) {
    if (!x || !y || !code || k < 1 || k > 3) return set_error(VU_ERR_INVALID, "maxpool_idx_fwd: bad args (k <= 3)");
    VU_TB_ALIGN((C | x_ctot | x_coff | y_ctot | y_coff) % 8, "maxpool_idx_fwd");
    const int Ho = (H + 2 * p - k) / s + 1, Wo = (W + 2 * p - k) / s + 1;
    maxpool_idx_fwd_kernel<<<tb_grid((size_t)N * Ho * Wo * (C / 8)), 256, 0, (cudaStream_t)st>>>(
        (const uint16_t*)x, x_ctot, x_coff, (uint16_t*)y, y_ctot, y_coff, (uint8_t*)code, N, H, W, C / 8, k, s, p, Ho, Wo, dtype);
    VU_CHECK_LAUNCH("maxpool_idx_fwd");
    return 0;
}

extern "C" int vu_maxpool_idx_bwd(const void* code, const void* dy, void* dx, int N, int H, int W, int C, int dy_ctot,
                                  int dy_coff, int dx_ctot, int dx_coff, int k, int s, int p, int dtype, vu_stream_t st) {
    if (!code || !dy || !dx || k < 1 || k > 3) return set_error(VU_ERR_INVALID, "maxpool_idx_bwd: bad args (k <= 3)");
    VU_TB_ALIGN((C | dy_ctot | dy_coff | dx_ctot | dx_coff) % 8, "maxpool_idx_bwd");
    const int Ho = (H + 2 * p - k) / s + 1, Wo = (W + 2 * p - k) / s + 1;
    if (k == 3 && s == 2 && p == 1 && C == 64 && H <= 65535 && N <= 65535) {
        maxpool_idx_bwd_k3s2_kernel<8><<<dim3((W * 8 + 255) / 256, H, N), 256, 0, (cudaStream_t)st>>>(
            (const uint8_t*)code, (const uint16_t*)dy, dy_ctot, dy_coff, (uint16_t*)dx, dx_ctot, dx_coff, H, W, Ho, Wo, dtype);
        VU_CHECK_LAUNCH("maxpool_idx_bwd_k3s2");
        return 0;
    }
    if ((size_t)N * H * W * (C / 8) < (1ull << 31))
        maxpool_idx_bwd_kernel<uint32_t><<<tb_grid((size_t)N * H * W * (C / 8)), 256, 0, (cudaStream_t)st>>>(
            (const uint8_t*)code, (const uint16_t*)dy, dy_ctot, dy_coff, (uint16_t*)dx, dx_ctot, dx_coff, N, H, W, C / 8, k, s, p,
            Ho, Wo, dtype);
    else
        maxpool_idx_bwd_kernel<size_t><<<tb_grid((size_t)N * H * W * (C / 8)), 256, 0, (cudaStream_t)st>>>(
            (const uint8_t*)code, (const uint16_t*)dy, dy_ctot, dy_coff, (uint16_t*)dx, dx_ctot, dx_coff, N, H, W, C / 8, k, s, p,
            Ho, Wo, dtype);
    VU_CHECK_LAUNCH("maxpool_idx_bwd");
    return 0;
}

extern "C" int vu_bilinear_ac_bwd(const void* dy, void* dx, int N, int Hs, int Ws, int C, int Ho, int Wo, int dy_ctot,
                                  int dy_coff, int dx_ctot, int dx_coff, int accumulate, int dtype, vu_stream_t st) {
    if (!dy || !dx) return set_error(VU_ERR_INVALID, "bilinear_bwd: null");
    VU_TB_ALIGN((C | dy_ctot | dy_coff | dx_ctot | dx_coff) % 8, "bilinear_bwd");
    const float sh = Ho > 1 ? (float)(Hs - 1) / (float)(Ho - 1) : 0.0f;
    const float sw = Wo > 1 ? (float)(Ws - 1) / (float)(Wo - 1) : 0.0f;
    // a source pixel is touched by at most 2/scale + 1 destination rows: the unrolled kernel covers the x2 upsamples
    static const bool generic_only = getenv("VU_BILINEAR_GENERIC") != nullptr;
    const bool taps_ok = sh >= 0.4f && sw >= 0.4f && (size_t)N * Hs * Ws * (C / 8) < (1ull << 31);
#define VU_BILB_CASE(c8)                                                                                                       \
    if (taps_ok && !generic_only && N <= 65535 && C / 8 == c8) {                                                               \
        launch_bilinear_bwd_run<c8>((const uint16_t*)dy, dy_ctot, dy_coff, (uint16_t*)dx, dx_ctot, dx_coff, N, Hs, Ws, Ho, Wo, sh, \
                                    sw, accumulate, dtype, (cudaStream_t)st);                                                  \
        VU_CHECK_LAUNCH("bilinear_bwd_run");                                                                                   \
        return 0;                                                                                                              \
    }
    VU_BILB_CASE(4) VU_BILB_CASE(8) VU_BILB_CASE(16) VU_BILB_CASE(32) VU_BILB_CASE(64)
#undef VU_BILB_CASE
    if (taps_ok)
        bilinear_bwd_taps_kernel<uint32_t><<<tb_grid((size_t)N * Hs * Ws * (C / 8)), 256, 0, (cudaStream_t)st>>>(
            (const uint16_t*)dy, dy_ctot, dy_coff, (uint16_t*)dx, dx_ctot, dx_coff, N, Hs, Ws, Ho, Wo, C / 8, sh, sw, accumulate,
            dtype);
    else
        bilinear_bwd_kernel<<<tb_grid((size_t)N * Hs * Ws * (C / 8)), 256, 0, (cudaStream_t)st>>>(
            (const uint16_t*)dy, dy_ctot, dy_coff, (uint16_t*)dx, dx_ctot, dx_coff, N, Hs, Ws, Ho, Wo, C / 8, sh, sw, accumulate,
            dtype);
    VU_CHECK_LAUNCH("bilinear_bwd");
    return 0;
}

extern "C" int vu_resize_map_bwd(const float* dy, float* dx, int N, int Hs, int Ws, int Ho, int Wo, vu_stream_t st) {
    if (!dy || !dx) return set_error(VU_ERR_INVALID, "resize_map_bwd: null");
    const float sh = Ho > 1 ? (float)(Hs - 1) / (float)(Ho - 1) : 0.0f;
    const float sw = Wo > 1 ? (float)(Ws - 1) / (float)(Wo - 1) : 0.0f;
    resize_map_bwd_kernel<<<tb_grid((size_t)N * Hs * Ws), 256, 0, (cudaStream_t)st>>>(dy, dx, N, Hs, Ws, Ho, Wo, sh, sw);
    VU_CHECK_LAUNCH("resize_map_bwd");
    return 0;
}

extern "C" int vu_scale_by_map(const void* a, int a_ctot, int a_coff, const float* m, void* out, int o_ctot, int o_coff,
                               size_t npix, int C, int accumulate, int dtype, vu_stream_t st) {
    if (!a || !m || !out) return set_error(VU_ERR_INVALID, "scale_by_map: null");
    VU_TB_ALIGN((C | a_ctot | a_coff | o_ctot | o_coff) % 8, "scale_by_map");
    scale_by_map_kernel<<<tb_grid(npix * (C / 8)), 256, 0, (cudaStream_t)st>>>((const uint16_t*)a, a_ctot, a_coff, m,
                                                                              (uint16_t*)out, o_ctot, o_coff, npix, C / 8,
                                                                              accumulate, dtype);
    VU_CHECK_LAUNCH("scale_by_map");
    return 0;
}

extern "C" int vu_channel_dot(const void* a, int a_ctot, int a_coff, const void* b, int b_ctot, int b_coff, float* out,
                              size_t npix, int C, int dtype, vu_stream_t st) {
    if (!a || !b || !out) return set_error(VU_ERR_INVALID, "channel_dot: null");
    VU_TB_ALIGN((C | a_ctot | a_coff | b_ctot | b_coff) % 8, "channel_dot");
    channel_dot_kernel<<<tb_grid(npix * 8), 256, 0, (cudaStream_t)st>>>((const uint16_t*)a, a_ctot, a_coff,
                                                                       (const uint16_t*)b, b_ctot, b_coff, out, npix, C / 8,
                                                                       dtype);
    VU_CHECK_LAUNCH("channel_dot");
    return 0;
}

extern "C" int vu_map_expand(const float* m, const float* w, const void* mask, int k_ctot, int k_coff, void* out,
                             int o_ctot, int o_coff, size_t npix, int C, int dtype, vu_stream_t st) {
    if (!m || !w || !out) return set_error(VU_ERR_INVALID, "map_expand: null");
    VU_TB_ALIGN((C | k_ctot | k_coff | o_ctot | o_coff) % 8, "map_expand");
    map_expand_kernel<<<tb_grid(npix * (C / 8)), 256, 0, (cudaStream_t)st>>>(m, w, (const uint16_t*)mask, k_ctot, k_coff,
                                                                            (uint16_t*)out, o_ctot, o_coff, npix, C / 8, dtype);
    VU_CHECK_LAUNCH("map_expand");
    return 0;
}

extern "C" int vu_weighted_channel_sum(const void* x, int x_ctot, int x_coff, const float* m, int n_img, size_t pix_per_img,
                                       int C, int per_image, double* sums, int dtype, vu_stream_t st) {
    if (!x || !sums || C < 8 || C > 1024 || n_img < 1) return set_error(VU_ERR_INVALID, "weighted_channel_sum: bad args");
    VU_TB_ALIGN((C | x_ctot | x_coff) % 8, "weighted_channel_sum");
    const int C8 = C / 8, lanes = 256 / C8;
    cudaStream_t s = (cudaStream_t)st;
    cudaMemsetAsync(sums, 0, (size_t)(per_image ? n_img : 1) * C * sizeof(double), s);
    int bpi = (int)((pix_per_img + (size_t)lanes * 16 - 1) / ((size_t)lanes * 16));
    const int cap = (num_sms() * 4 + n_img - 1) / n_img;
    if (bpi > cap) bpi = cap;
    if (bpi < 1) bpi = 1;
    weighted_channel_sum_kernel<<<n_img * bpi, 256, lanes * C * sizeof(float), s>>>((const uint16_t*)x, x_ctot, x_coff, m,
                                                                                  pix_per_img, C8, per_image, bpi, sums, dtype);
    VU_CHECK_LAUNCH("weighted_channel_sum");
    return 0;
}
