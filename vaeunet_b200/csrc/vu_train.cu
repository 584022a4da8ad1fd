// Memory-bound kernels of the training step (train.py:385-411): train-mode BatchNorm2d forward /
// backward over NHWC 16-bit maps, gradient accumulation, and the backward passes of max-pool, bilinear
// resize, attention gate, output conv and the small latent-vector layers. All arithmetic is fp32 (fp64 for the
// cross-block accumulation of batch statistics); activations and activation gradients are 16-bit NHWC slices.
#include <cstdlib>
#include "vu_half.cuh"
#include "vu_internal.h"

namespace vu {

// Tuning knobs of the BatchNorm launches (A/B runs): CTAs per SM of the statistics / apply / backward grids. One resident
// wave each (2 / 3 / 2 CTAs of 256 threads per SM at 118 / 78 / 122 registers): with 8-16 CTAs per SM the per-CTA prologue
// (coefficients from the fp64 sums -> smem -> sync) was paid 3-4 times in sequence on the 4-70 MB maps of this network
// (in-situ CUPTI sums per step, profiles/r2v_bn_grid.log: apply 0.69 -> 0.59 ms, backward apply 1.00 -> 0.82 ms).
static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v ? atoi(v) : dflt;
}
static int bn_mult_sums() { static const int m = env_int("VU_BN_MULT_SUMS", 2); return m; }
static int bn_mult_apply() { static const int m = env_int("VU_BN_MULT_APPLY", 3); return m; }
static int bn_mult_bwd() { static const int m = env_int("VU_BN_MULT_BWD", 2); return m; }

static int tr_grid(size_t items, int per_block, int mult = 8) {
    size_t g = (items + per_block - 1) / per_block;
    const size_t cap = (size_t)num_sms() * mult;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

// ---------------------------------------------------------------- per-channel sums over pixels
// Block = 256 threads = (256 / C8) pixel lanes x C8 channel chunks (C8 = C/8 <= 128). Each thread accumulates
// its 8 channels over a strided set of pixels in fp32, lanes are combined through smem, and one thread per
// channel adds the block's partial to the global fp64 accumulator.
// MODE 0: sums[c] += x, sums[C+c] += x^2                                    (BN forward statistics)
// MODE 1: dz = dy * (y > 0 if y), xh = (x - mean) * invstd: sums[c] += dz, sums[C+c] += dz * xh   (BN backward)
template <int MODE>
__global__ void __launch_bounds__(256, 2)
channel_sums_kernel(const uint16_t* __restrict__ x, int x_ctot, int x_coff, const uint16_t* __restrict__ dy,
                    int dy_ctot, int dy_coff, const uint16_t* __restrict__ y, int y_ctot, int y_coff,
                    const float* __restrict__ mean, const float* __restrict__ invstd, size_t npix, int C8,
                    double* __restrict__ sums, int dtype, const float* __restrict__ gamma = nullptr,
                    const float* __restrict__ relu_beta = nullptr) {
    extern __shared__ float sm[];   // [2][lanes][C8*8]
    const int lanes = 256 / C8;
    const int c8 = threadIdx.x % C8, pl = threadIdx.x / C8;
    const int C = C8 * 8;
    float a0[8], a1[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { a0[j] = 0.f; a1[j] = 0.f; }
    float mu[8], msc[8], msh[8];
    if (MODE == 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            mu[j] = mean[c8 * 8 + j];
            // ReLU mask recomputed from x exactly as bn_apply_kernel computed the output (no read of y)
            msc[j] = relu_beta ? invstd[c8 * 8 + j] * gamma[c8 * 8 + j] : 0.f;
            msh[j] = relu_beta ? relu_beta[c8 * 8 + j] - mu[j] * msc[j] : 0.f;
        }
    }
    if (pl < lanes) {
        const size_t step = (size_t)gridDim.x * lanes;
        size_t p = (size_t)blockIdx.x * lanes + pl;
        if (MODE == 0) {
            // four independent 16-byte loads in flight per thread
            for (; p + 3 * step < npix; p += 4 * step) {
                uint4 v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) v[u] = __ldg(reinterpret_cast<const uint4*>(x + (p + u * step) * x_ctot + x_coff) + c8);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    float f[8];
                    unpack8(v[u], f, dtype);
#pragma unroll
                    for (int j = 0; j < 8; ++j) { a0[j] += f[j]; a1[j] = fmaf(f[j], f[j], a1[j]); }
                }
            }
            for (; p < npix; p += step) {
                float f[8];
                unpack8(__ldg(reinterpret_cast<const uint4*>(x + p * x_ctot + x_coff) + c8), f, dtype);
#pragma unroll
                for (int j = 0; j < 8; ++j) { a0[j] += f[j]; a1[j] = fmaf(f[j], f[j], a1[j]); }
            }
        } else {
            // a1 accumulates dz * (x - mean); invstd is applied once per block below (8 registers less in the loop)
            auto body = [&](const uint4& vx, const uint4& vg, const uint4& vy) {
                float f[8], g[8];
                unpack8(vx, f, dtype);
                unpack8(vg, g, dtype);
                if (y != nullptr) {
                    float o[8];
                    unpack8(vy, o, dtype);
#pragma unroll
                    for (int j = 0; j < 8; ++j) g[j] = o[j] > 0.f ? g[j] : 0.f;
                } else if (relu_beta != nullptr) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) g[j] = fmaf(f[j], msc[j], msh[j]) > 0.f ? g[j] : 0.f;
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) { a0[j] += g[j]; a1[j] = fmaf(g[j], f[j] - mu[j], a1[j]); }
            };
            constexpr int U = 4;                              // pixels in flight per thread: 8-12 independent 16-byte loads
            for (; p + (U - 1) * step < npix; p += U * step) {
                uint4 vx[U], vg[U], vy[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const size_t q = p + u * step;
                    vx[u] = __ldg(reinterpret_cast<const uint4*>(x + q * x_ctot + x_coff) + c8);
                    vg[u] = __ldg(reinterpret_cast<const uint4*>(dy + q * dy_ctot + dy_coff) + c8);
                    vy[u] = y ? __ldg(reinterpret_cast<const uint4*>(y + q * y_ctot + y_coff) + c8) : make_uint4(0, 0, 0, 0);
                }
#pragma unroll
                for (int u = 0; u < U; ++u) body(vx[u], vg[u], vy[u]);
            }
            for (; p < npix; p += step) {
                const uint4 vx = __ldg(reinterpret_cast<const uint4*>(x + p * x_ctot + x_coff) + c8);
                const uint4 vg = __ldg(reinterpret_cast<const uint4*>(dy + p * dy_ctot + dy_coff) + c8);
                const uint4 vy = y ? __ldg(reinterpret_cast<const uint4*>(y + p * y_ctot + y_coff) + c8) : make_uint4(0, 0, 0, 0);
                body(vx, vg, vy);
            }
            if (MODE == 1) {
#pragma unroll
                for (int j = 0; j < 8; ++j) a1[j] *= invstd[c8 * 8 + j];
            }
        }
    }
    float* s0 = sm;
    float* s1 = sm + lanes * C;
    if (pl < lanes) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { s0[pl * C + c8 * 8 + j] = a0[j]; s1[pl * C + c8 * 8 + j] = a1[j]; }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += 256) {
        float t0 = 0.f, t1 = 0.f;
        for (int l = 0; l < lanes; ++l) { t0 += s0[l * C + c]; t1 += s1[l * C + c]; }
        atomicAdd(&sums[c], (double)t0);
        atomicAdd(&sums[C + c], (double)t1);
    }
}

// mean / invstd of channel c from the fp64 sums (same expressions in every kernel that derives them, so the forward's
// coefficients and the statistics saved for the backward agree bit for bit). fp64 multiplies only: an fp64 division or
// square root per thread made these prologues longer than the loops they precede.
__device__ __forceinline__ void bn_stat_from_sums(const double* __restrict__ sums, int C, int c, double inv_count, float eps,
                                                  float& mean, float& invstd, double& var_out) {
    const double m = sums[c] * inv_count;
    double var = sums[C + c] * inv_count - m * m;
    if (var < 0.0) var = 0.0;
    mean = (float)m;
    invstd = 1.0f / sqrtf((float)var + eps);
    var_out = var;
}

// mean / invstd from the fp64 sums; running statistics updated like nn.BatchNorm2d (momentum, unbiased var).
__global__ void bn_finalize_kernel(const double* __restrict__ sums, double count, int C, float eps, float momentum,
                                   float* __restrict__ mean, float* __restrict__ invstd,
                                   float* __restrict__ running_mean, float* __restrict__ running_var) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    float mf, isf; double var;
    bn_stat_from_sums(sums, C, c, 1.0 / count, eps, mf, isf, var);
    const double m = (double)mf;
    mean[c] = mf;
    invstd[c] = isf;
    if (running_mean != nullptr) {
        const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)m;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
    }
}

// Running statistics of nn.BatchNorm2d from the saved batch statistics (the functional torch custom ops update them in a
// separate, mutating op): var = 1 / invstd^2 - eps (biased), running_var takes the unbiased estimate.
__global__ void bn_update_running_kernel(const float* __restrict__ mean, const float* __restrict__ invstd, double count,
                                         float eps, float momentum, float* __restrict__ running_mean,
                                         float* __restrict__ running_var, int C) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const double is = (double)invstd[c];
    double var = 1.0 / (is * is) - (double)eps;
    if (var < 0.0) var = 0.0;
    const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean[c];
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
}

// y = act(gamma * (x - mean) * invstd + beta + residual) = act(x * sc + sh + residual).
// The launch uses gridDim * 256 threads with 256 % C8 == 0 whenever C8 is a power of two (every BatchNorm of the model),
// so a thread keeps the same 8 channels for its whole grid-stride loop and loads the coefficients once (FIXED = true).
// SUMS variant (sums != nullptr): the batch statistics are derived in the kernel from the fp64 sums the producing conv (or
// vu_bn_stats) accumulated - no separate finalize launch; block 0 also writes mean / invstd for the backward and updates the
// running statistics like nn.BatchNorm2d (momentum, unbiased variance).
template <bool FIXED>
__global__ void __launch_bounds__(256)
bn_apply_kernel(const uint16_t* __restrict__ x, int x_ctot, int x_coff, const float* __restrict__ mean,
                const float* __restrict__ invstd, const float* __restrict__ gamma, const float* __restrict__ beta,
                const uint16_t* __restrict__ res, int res_ctot, int res_coff, int relu, uint16_t* __restrict__ y,
                int y_ctot, int y_coff, size_t npix, int C8, int dtype, const double* __restrict__ sums = nullptr,
                double count = 1.0, double inv_count = 1.0, float eps = 0.f, float momentum = 0.f,
                float* __restrict__ running_mean = nullptr,
                float* __restrict__ running_var = nullptr, float* __restrict__ mean_out = nullptr,
                float* __restrict__ invstd_out = nullptr) {
    const size_t total = npix * C8;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const int C = C8 * 8;
    if (sums != nullptr && blockIdx.x == 0) {
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            float m, is; double var;
            bn_stat_from_sums(sums, C, c, inv_count, eps, m, is, var);
            mean_out[c] = m; invstd_out[c] = is;
            if (running_mean != nullptr) {
                const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
                running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m;
                running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
            }
        }
    }
    // y = x * sc + sh: the per-channel coefficients are computed ONCE per block into shared memory (C / 256 channels per
    // thread) - with a per-thread prologue over its 8 channels the fp64 -> fp32 statistics arithmetic cost more than the
    // 4..14 chunks a thread then processes
    extern __shared__ float s_coef[];
    float* s_sc = s_coef;
    float* s_sh = s_coef + C;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float m, is;
        if (sums != nullptr) { double var; bn_stat_from_sums(sums, C, c, inv_count, eps, m, is, var); }
        else { m = mean[c]; is = invstd[c]; }
        const float a = is * gamma[c];
        s_sc[c] = a;
        s_sh[c] = beta[c] - m * a;
    }
    __syncthreads();
    float sc[8], sh[8];
    auto coeff = [&](int c8) {
        const float4 a0 = *reinterpret_cast<const float4*>(s_sc + c8 * 8), a1 = *reinterpret_cast<const float4*>(s_sc + c8 * 8 + 4);
        const float4 b0 = *reinterpret_cast<const float4*>(s_sh + c8 * 8), b1 = *reinterpret_cast<const float4*>(s_sh + c8 * 8 + 4);
        sc[0] = a0.x; sc[1] = a0.y; sc[2] = a0.z; sc[3] = a0.w; sc[4] = a1.x; sc[5] = a1.y; sc[6] = a1.z; sc[7] = a1.w;
        sh[0] = b0.x; sh[1] = b0.y; sh[2] = b0.z; sh[3] = b0.w; sh[4] = b1.x; sh[5] = b1.y; sh[6] = b1.z; sh[7] = b1.w;
    };
    size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    auto one = [&](size_t p, int c8, const uint4& vx, const uint4& vr) {
        if (!FIXED) coeff(c8);
        float f[8], r[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        unpack8(vx, f, dtype);
        if (res != nullptr) unpack8(vr, r, dtype);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float v = fmaf(f[j], sc[j], sh[j]) + r[j];
            f[j] = relu ? fmaxf(v, 0.f) : v;
        }
        *(reinterpret_cast<uint4*>(y + p * y_ctot + y_coff) + c8) = pack8(f, dtype);
    };
    auto ldx = [&](size_t p, int c8) { return __ldg(reinterpret_cast<const uint4*>(x + p * x_ctot + x_coff) + c8); };
    auto ldr = [&](size_t p, int c8) {
        return res ? __ldg(reinterpret_cast<const uint4*>(res + p * res_ctot + res_coff) + c8) : make_uint4(0, 0, 0, 0);
    };
    if (FIXED) {
        // 256 % C8 == 0: the thread's channel group never changes and its pixel index advances by a constant; no
        // division in the loop (64-bit div/mod per element made these kernels instruction-bound, not HBM-bound)
        const int c8 = threadIdx.x % C8;
        const size_t pstep = stride / C8;
        size_t p = idx / C8;
        coeff(c8);
        for (; p + 3 * pstep < npix; p += 4 * pstep) {
            uint4 vx[4], vr[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { vx[u] = ldx(p + u * pstep, c8); vr[u] = ldr(p + u * pstep, c8); }
#pragma unroll
            for (int u = 0; u < 4; ++u) one(p + u * pstep, c8, vx[u], vr[u]);
        }
        for (; p < npix; p += pstep) one(p, c8, ldx(p, c8), ldr(p, c8));
    } else {
        for (; idx < total; idx += stride) {
            const size_t p = idx / C8;
            const int c8 = (int)(idx % C8);
            one(p, c8, ldx(p, c8), ldr(p, c8));
        }
    }
}

// dx = gamma * invstd * (dz - mean(dz) - xh * mean(dz * xh)) = ka * dz + kb * x + kc,  dz = dy * (y > 0 if y).
// dz_out (optional) = dz (the gradient of a residual branch added before the ReLU). Coefficients per thread as above.
template <bool FIXED>
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const uint16_t* __restrict__ x, int x_ctot, int x_coff, const uint16_t* __restrict__ dy,
                    int dy_ctot, int dy_coff, const uint16_t* __restrict__ y, int y_ctot, int y_coff,
                    const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ gamma,
                    const double* __restrict__ sums, double count, uint16_t* __restrict__ dx, int dx_ctot, int dx_coff,
                    uint16_t* __restrict__ dz_out, int dz_ctot, int dz_coff, size_t npix, int C8, int dtype,
                    float* __restrict__ dgamma, float* __restrict__ dbeta, const float* __restrict__ relu_beta) {
    const size_t total = npix * C8;
    const int C = C8 * 8;
    if (blockIdx.x == 0 && dgamma != nullptr)      // parameter gradients, accumulated (gradient accumulation steps)
        for (int c = threadIdx.x; c < C; c += blockDim.x) { dgamma[c] += (float)sums[C + c]; dbeta[c] += (float)sums[c]; }
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const float inv_count = (float)(1.0 / count);
    // coefficients once per block in shared memory (see bn_apply_kernel): dx = ka * dz + kb * x + kc, mask = x * msc + msh > 0
    extern __shared__ float s_coef[];
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        // fp32 from here on: fp64 division is ~100x slower than an fp32 multiply
        const float m_dz = (float)sums[c] * inv_count, m_dzx = (float)sums[C + c] * inv_count;
        const float is = invstd[c], a = gamma[c] * is;
        const float b = -a * m_dzx * is;
        s_coef[c] = a;
        s_coef[C + c] = b;
        s_coef[2 * C + c] = -a * m_dz - b * mean[c];
        s_coef[3 * C + c] = is * gamma[c];                                   // same expressions as bn_apply_kernel
        s_coef[4 * C + c] = relu_beta ? relu_beta[c] - mean[c] * (is * gamma[c]) : 0.f;
    }
    __syncthreads();
    float ka[8], kb[8], kc[8], msc[8], msh[8];
    auto coeff = [&](int c8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int c = c8 * 8 + j;
            ka[j] = s_coef[c]; kb[j] = s_coef[C + c]; kc[j] = s_coef[2 * C + c];
            msc[j] = s_coef[3 * C + c]; msh[j] = s_coef[4 * C + c];
        }
    };
    size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    auto one = [&](size_t p, int c8, const uint4& vx, const uint4& vg, const uint4& vy) {
        if (!FIXED) coeff(c8);
        float f[8], g[8];
        unpack8(vx, f, dtype);
        unpack8(vg, g, dtype);
        if (y != nullptr) {
            float o[8];
            unpack8(vy, o, dtype);
#pragma unroll
            for (int j = 0; j < 8; ++j) g[j] = o[j] > 0.f ? g[j] : 0.f;
        } else if (relu_beta != nullptr) {
#pragma unroll
            for (int j = 0; j < 8; ++j) g[j] = fmaf(f[j], msc[j], msh[j]) > 0.f ? g[j] : 0.f;
        }
        if (dz_out != nullptr) *(reinterpret_cast<uint4*>(dz_out + p * dz_ctot + dz_coff) + c8) = pack8(g, dtype);
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = fmaf(ka[j], g[j], fmaf(kb[j], f[j], kc[j]));
        *(reinterpret_cast<uint4*>(dx + p * dx_ctot + dx_coff) + c8) = pack8(o, dtype);
    };
    auto ld = [&](const uint16_t* t, int ctot, int coff, size_t p, int c8) {
        return __ldg(reinterpret_cast<const uint4*>(t + p * ctot + coff) + c8);
    };
    if (FIXED) {     // fixed channel group, constant pixel step: no division in the loop (see bn_apply_kernel)
        const int c8 = threadIdx.x % C8;
        const size_t pstep = stride / C8;
        size_t p = idx / C8;
        coeff(c8);
        for (; p + 3 * pstep < npix; p += 4 * pstep) {       // 8..12 independent 16-byte loads in flight per thread
            uint4 vx[4], vg[4], vy[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const size_t q = p + u * pstep;
                vx[u] = ld(x, x_ctot, x_coff, q, c8);
                vg[u] = ld(dy, dy_ctot, dy_coff, q, c8);
                vy[u] = y ? ld(y, y_ctot, y_coff, q, c8) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) one(p + u * pstep, c8, vx[u], vg[u], vy[u]);
        }
        for (; p < npix; p += pstep)
            one(p, c8, ld(x, x_ctot, x_coff, p, c8), ld(dy, dy_ctot, dy_coff, p, c8),
                y ? ld(y, y_ctot, y_coff, p, c8) : make_uint4(0, 0, 0, 0));
    } else {
        for (; idx < total; idx += stride) {
            const size_t p = idx / C8;
            const int c8 = (int)(idx % C8);
            one(p, c8, ld(x, x_ctot, x_coff, p, c8), ld(dy, dy_ctot, dy_coff, p, c8),
                y ? ld(y, y_ctot, y_coff, p, c8) : make_uint4(0, 0, 0, 0));
        }
    }
}

// dst += src (gradient accumulation over the consumers of an activation)
__global__ void __launch_bounds__(256)
add_nhwc_kernel(uint16_t* __restrict__ dst, int d_ctot, int d_coff, const uint16_t* __restrict__ src, int s_ctot,
                int s_coff, size_t npix, int C8, int dtype) {
    const size_t total = npix * C8;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int c8 = idx % C8;
        const size_t p = idx / C8;
        uint4* dp = reinterpret_cast<uint4*>(dst + p * d_ctot + d_coff) + c8;
        float a[8], b[8];
        unpack8(*dp, a, dtype);
        unpack8(__ldg(reinterpret_cast<const uint4*>(src + p * s_ctot + s_coff) + c8), b, dtype);
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] += b[j];
        *dp = pack8(a, dtype);
    }
}

}  // namespace vu

using namespace vu;

#define VU_TR_ALIGN(cond, what) if (cond) return set_error(VU_ERR_ALIGN, what ": channel counts/offsets must be multiples of 8")

extern "C" int vu_bn_stats(const void* x, size_t npix, int C, int x_ctot, int x_coff, double* sums, int dtype, vu_stream_t st) {
    if (!x || !sums || C < 8 || C > 1024) return set_error(VU_ERR_INVALID, "bn_stats: bad args (8 <= C <= 1024)");
    VU_TR_ALIGN((C | x_ctot | x_coff) % 8, "bn_stats");
    const int C8 = C / 8, lanes = 256 / C8;
    cudaMemsetAsync(sums, 0, 2 * C * sizeof(double), (cudaStream_t)st);
    channel_sums_kernel<0><<<tr_grid(npix, lanes * 16, 8), 256, 2 * lanes * C * sizeof(float), (cudaStream_t)st>>>(
        (const uint16_t*)x, x_ctot, x_coff, nullptr, 0, 0, nullptr, 0, 0, nullptr, nullptr, npix, C8, sums, dtype);
    VU_CHECK_LAUNCH("bn_stats");
    return 0;
}

extern "C" int vu_bn_finalize(const double* sums, double count, int C, float eps, float momentum, float* mean,
                              float* invstd, float* running_mean, float* running_var, vu_stream_t st) {
    if (!sums || !mean || !invstd || C < 1 || count < 1) return set_error(VU_ERR_INVALID, "bn_finalize: bad args");
    bn_finalize_kernel<<<(C + 127) / 128, 128, 0, (cudaStream_t)st>>>(sums, count, C, eps, momentum, mean, invstd,
                                                                       running_mean, running_var);
    VU_CHECK_LAUNCH("bn_finalize");
    return 0;
}

extern "C" int vu_bn_update_running(const float* mean, const float* invstd, double count, int C, float eps, float momentum,
                                    float* running_mean, float* running_var, vu_stream_t st) {
    if (!mean || !invstd || !running_mean || !running_var || C < 1 || count < 1) return set_error(VU_ERR_INVALID, "bn_update_running: bad args");
    bn_update_running_kernel<<<(C + 127) / 128, 128, 0, (cudaStream_t)st>>>(mean, invstd, count, eps, momentum, running_mean,
                                                                            running_var, C);
    VU_CHECK_LAUNCH("bn_update_running");
    return 0;
}

extern "C" int vu_bn_apply(const void* x, size_t npix, int C, int x_ctot, int x_coff, const float* mean,
                           const float* invstd, const float* gamma, const float* beta, const void* residual,
                           int res_ctot, int res_coff, int relu, void* y, int y_ctot, int y_coff, int dtype,
                           vu_stream_t st) {
    if (!x || !y || !mean || !invstd || !gamma || !beta) return set_error(VU_ERR_INVALID, "bn_apply: null");
    VU_TR_ALIGN((C | x_ctot | x_coff | y_ctot | y_coff | res_ctot | res_coff) % 8, "bn_apply");
    const int C8 = C / 8;
    const int grid = tr_grid(npix * C8, 256 * 4, bn_mult_apply());
    if (256 % C8 == 0)
        bn_apply_kernel<true><<<grid, 256, 2 * C * sizeof(float), (cudaStream_t)st>>>(
            (const uint16_t*)x, x_ctot, x_coff, mean, invstd, gamma, beta, (const uint16_t*)residual, res_ctot, res_coff, relu,
            (uint16_t*)y, y_ctot, y_coff, npix, C8, dtype);
    else
        bn_apply_kernel<false><<<grid, 256, 2 * C * sizeof(float), (cudaStream_t)st>>>(
            (const uint16_t*)x, x_ctot, x_coff, mean, invstd, gamma, beta, (const uint16_t*)residual, res_ctot, res_coff, relu,
            (uint16_t*)y, y_ctot, y_coff, npix, C8, dtype);
    VU_CHECK_LAUNCH("bn_apply");
    return 0;
}

extern "C" int vu_bn_apply_sums(const void* x, size_t npix, int C, int x_ctot, int x_coff, const double* sums, float eps,
                                float momentum, float* running_mean, float* running_var, float* mean, float* invstd,
                                const float* gamma, const float* beta, const void* residual, int res_ctot, int res_coff,
                                int relu, void* y, int y_ctot, int y_coff, int dtype, vu_stream_t st) {
    if (!x || !y || !sums || !mean || !invstd || !gamma || !beta || npix < 1) return set_error(VU_ERR_INVALID, "bn_apply_sums: null");
    VU_TR_ALIGN((C | x_ctot | x_coff | y_ctot | y_coff | res_ctot | res_coff) % 8, "bn_apply_sums");
    const int C8 = C / 8;
    const int grid = tr_grid(npix * C8, 256 * 4, bn_mult_apply());
    if (256 % C8 == 0)
        bn_apply_kernel<true><<<grid, 256, 2 * C * sizeof(float), (cudaStream_t)st>>>(
            (const uint16_t*)x, x_ctot, x_coff, nullptr, nullptr, gamma, beta, (const uint16_t*)residual, res_ctot, res_coff, relu,
            (uint16_t*)y, y_ctot, y_coff, npix, C8, dtype, sums, (double)npix, 1.0 / (double)npix, eps, momentum, running_mean,
            running_var, mean, invstd);
    else
        bn_apply_kernel<false><<<grid, 256, 2 * C * sizeof(float), (cudaStream_t)st>>>(
            (const uint16_t*)x, x_ctot, x_coff, nullptr, nullptr, gamma, beta, (const uint16_t*)residual, res_ctot, res_coff, relu,
            (uint16_t*)y, y_ctot, y_coff, npix, C8, dtype, sums, (double)npix, 1.0 / (double)npix, eps, momentum, running_mean,
            running_var, mean, invstd);
    VU_CHECK_LAUNCH("bn_apply_sums");
    return 0;
}

extern "C" int vu_bn_bwd(const void* x, size_t npix, int C, int x_ctot, int x_coff, const void* dy, int dy_ctot,
                         int dy_coff, const void* y_relu, int y_ctot, int y_coff, const float* mean,
                         const float* invstd, const float* gamma, double* sums, void* dx, int dx_ctot, int dx_coff,
                         void* dz_out, int dz_ctot, int dz_coff, float* dgamma, float* dbeta, const float* relu_beta, int dtype,
                         vu_stream_t st) {
    if (!x || !dy || !mean || !invstd || !gamma || !sums || !dx || C < 8 || C > 1024 || (!dgamma != !dbeta))
        return set_error(VU_ERR_INVALID, "bn_bwd: bad args");
    VU_TR_ALIGN((C | x_ctot | x_coff | dy_ctot | dy_coff | y_ctot | y_coff | dx_ctot | dx_coff | dz_ctot | dz_coff) % 8, "bn_bwd");
    const int C8 = C / 8, lanes = 256 / C8;
    cudaStream_t s = (cudaStream_t)st;
    cudaMemsetAsync(sums, 0, 2 * C * sizeof(double), s);
    channel_sums_kernel<1><<<tr_grid(npix, lanes * 8, bn_mult_sums()), 256, 2 * lanes * C * sizeof(float), s>>>(
        (const uint16_t*)x, x_ctot, x_coff, (const uint16_t*)dy, dy_ctot, dy_coff, (const uint16_t*)y_relu, y_ctot, y_coff,
        mean, invstd, npix, C8, sums, dtype, gamma, y_relu ? nullptr : relu_beta);
    VU_CHECK_LAUNCH("bn_bwd_reduce");
    const int grid = tr_grid(npix * C8, 256 * 4, bn_mult_bwd());
    if (256 % C8 == 0)
        bn_bwd_apply_kernel<true><<<grid, 256, 5 * C * sizeof(float), s>>>(
            (const uint16_t*)x, x_ctot, x_coff, (const uint16_t*)dy, dy_ctot, dy_coff, (const uint16_t*)y_relu, y_ctot, y_coff,
            mean, invstd, gamma, sums, (double)npix, (uint16_t*)dx, dx_ctot, dx_coff, (uint16_t*)dz_out, dz_ctot, dz_coff, npix,
            C8, dtype, dgamma, dbeta, y_relu ? nullptr : relu_beta);
    else
        bn_bwd_apply_kernel<false><<<grid, 256, 5 * C * sizeof(float), s>>>(
            (const uint16_t*)x, x_ctot, x_coff, (const uint16_t*)dy, dy_ctot, dy_coff, (const uint16_t*)y_relu, y_ctot, y_coff,
            mean, invstd, gamma, sums, (double)npix, (uint16_t*)dx, dx_ctot, dx_coff, (uint16_t*)dz_out, dz_ctot, dz_coff, npix,
            C8, dtype, dgamma, dbeta, y_relu ? nullptr : relu_beta);
    VU_CHECK_LAUNCH("bn_bwd_apply");
    return 0;
}

extern "C" int vu_add_nhwc(void* dst, int d_ctot, int d_coff, const void* src, int s_ctot, int s_coff, size_t npix,
                           int C, int dtype, vu_stream_t st) {
    if (!dst || !src) return set_error(VU_ERR_INVALID, "add_nhwc: null");
    VU_TR_ALIGN((C | d_ctot | d_coff | s_ctot | s_coff) % 8, "add_nhwc");
    add_nhwc_kernel<<<tr_grid(npix * (C / 8), 256), 256, 0, (cudaStream_t)st>>>((uint16_t*)dst, d_ctot, d_coff,
                                                                               (const uint16_t*)src, s_ctot, s_coff, npix,
                                                                               C / 8, dtype);
    VU_CHECK_LAUNCH("add_nhwc");
    return 0;
}
