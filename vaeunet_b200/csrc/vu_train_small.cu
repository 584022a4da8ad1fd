// Small fp32 pieces of the training step: the single-channel BatchNorm + sigmoid of the attention
// gate's psi branch (unet_resnet.py:17-21) forward/backward, the dense layers that act on spatially constant maps
// or pooled vectors (z_initial, z_proj, mu/logvar heads: 1x1 conv [+ train-mode BN] [+ ReLU] == per-image mat-vec),
// the reparameterisation backward, the stem 7x7 weight gradient, and gradient-norm + AdamW (train.py:403-408).
#include "vu_half.cuh"
#include "vu_internal.h"

namespace vu {

// ---------------------------------------------------------------- fp32 map reductions
// MODE 0: s0 += x, s1 += x^2.   MODE 1 (psi backward): dp = dpsi * psi * (1 - psi), ph = (x - mean) * invstd:
// s0 += dp, s1 += dp * ph.
template <int MODE>
__global__ void __launch_bounds__(256)
map_sums_kernel(const float* __restrict__ x, const float* __restrict__ dpsi, const float* __restrict__ psi,
                const float* __restrict__ stat, size_t n, double* __restrict__ sums) {
    float a0 = 0.f, a1 = 0.f;
    const float mean = MODE ? stat[0] : 0.f, invstd = MODE ? stat[1] : 0.f;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float v = x[i];
        if (MODE == 0) { a0 += v; a1 = fmaf(v, v, a1); }
        else {
            const float s = psi[i], dp = dpsi[i] * s * (1.f - s);
            a0 += dp; a1 = fmaf(dp, (v - mean) * invstd, a1);
        }
    }
    a0 = warp_sum(a0); a1 = warp_sum(a1);
    __shared__ float sh[2][8];
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = a0; sh[1][threadIdx.x >> 5] = a1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float t0 = 0.f, t1 = 0.f;
        for (int w = 0; w < 8; ++w) { t0 += sh[0][w]; t1 += sh[1][w]; }
        atomicAdd(&sums[0], (double)t0);
        atomicAdd(&sums[1], (double)t1);
    }
}

// stat[0] = mean, stat[1] = invstd (+ running stats), from sums
__global__ void map_bn_finalize_kernel(const double* sums, double count, float eps, float momentum, float* stat,
                                       float* running_mean, float* running_var) {
    const double m = sums[0] / count;
    double var = sums[1] / count - m * m;
    if (var < 0.0) var = 0.0;
    stat[0] = (float)m;
    stat[1] = (float)(1.0 / sqrt(var + (double)eps));
    if (running_mean) {
        const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
        running_mean[0] = (1.f - momentum) * running_mean[0] + momentum * (float)m;
        running_var[0] = (1.f - momentum) * running_var[0] + momentum * (float)unbiased;
    }
}

__global__ void map_bn_sigmoid_apply_kernel(const float* __restrict__ x, const float* __restrict__ stat,
                                            const float* __restrict__ gamma_p, const float* __restrict__ beta_p,
                                            float* __restrict__ psi, size_t n) {
    const float mean = stat[0], invstd = stat[1], gamma = gamma_p[0], beta = beta_p[0];
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        psi[i] = 1.0f / (1.0f + expf(-((x[i] - mean) * invstd * gamma + beta)));
}

// dx = gamma * invstd * (dp - mean(dp) - ph * mean(dp * ph));  grads[0] = dgamma = sum dp*ph, grads[1] = dbeta = sum dp
__global__ void map_bn_sigmoid_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dpsi,
                                          const float* __restrict__ psi, const float* __restrict__ stat,
                                          const float* __restrict__ gamma_p,
                                          const double* __restrict__ sums, double count, float* __restrict__ dx,
                                          float* __restrict__ dgamma, float* __restrict__ dbeta, size_t n) {
    const float mean = stat[0], invstd = stat[1], gamma = gamma_p[0];
    const float m0 = (float)(sums[0] / count), m1 = (float)(sums[1] / count);
    if (blockIdx.x == 0 && threadIdx.x == 0) { dgamma[0] += (float)sums[1]; dbeta[0] += (float)sums[0]; }
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float s = psi[i], dp = dpsi[i] * s * (1.f - s);
        dx[i] = gamma * invstd * (dp - m0 - (x[i] - mean) * invstd * m1);
    }
}

// ---------------------------------------------------------------- dense layers on vectors
//   y_raw[m,c] = b[c] + sum_l z[m,l] w[c,l];  y = act(gamma (y_raw - mean) invstd + beta)
// One WARP per output channel (8 channels per block): the lanes split the L inputs of each dot product (coalesced rows of z
// and w) and combine with a shuffle butterfly; afterwards lane m % 32 holds row m of the channel, so the train-mode BatchNorm
// statistics over the M rows (every pixel of a spatially constant map is the same vector) are warp reductions too. All
// orders are fixed: results are bit-reproducible from run to run. (The first version gave a thread whole dot products -
// strided weight reads on 4..16 blocks: 148 us for the 512 -> 128 gate layer; this one is a few microseconds.)
constexpr int kVecM = 64;
__global__ void __launch_bounds__(256)
vec_layer_fwd_kernel(const float* __restrict__ z, const float* __restrict__ w, const float* __restrict__ b, int M, int L,
                     int C, const float* __restrict__ gamma, const float* __restrict__ beta, float eps, float momentum,
                     double stat_count, float* running_mean, float* running_var, int relu, float* __restrict__ y_raw,
                     float* __restrict__ y, float* __restrict__ mean_o, float* __restrict__ invstd_o) {
    const int lane = threadIdx.x & 31, c = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (c >= C) return;                                   // warp-uniform
    const float* wr = w + (size_t)c * L;
    const float bias = b ? b[c] : 0.f;
    float v[2] = {bias, bias};                            // rows lane, lane + 32
    // 8 columns per lane per pass, the filter row in registers; every load is unconditional (clamped index, zero weight for
    // the columns beyond L) so the 8 loads of a row are in flight together - predicated loads compile to branches that
    // serialise on the L2 latency
    for (int l0 = 0; l0 < L; l0 += 256) {
        float wv[8]; int li[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int l = l0 + lane + 32 * u;
            li[u] = min(l, L - 1);
            const float t = __ldg(wr + li[u]);
            wv[u] = l < L ? t : 0.f;
        }
        for (int m = 0; m < M; ++m) {
            const float* zr = z + (size_t)m * L;
            float acc = 0.f;
#pragma unroll
            for (int u = 0; u < 8; ++u) acc = fmaf(__ldg(zr + li[u]), wv[u], acc);
            acc = warp_sum(acc);
            if ((m & 31) == lane) { if (m < 32) v[0] += acc; else v[1] += acc; }
        }
    }
    const bool has0 = lane < M, has1 = lane + 32 < M;
    if (has0) y_raw[(size_t)lane * C + c] = v[0];
    if (has1) y_raw[(size_t)(lane + 32) * C + c] = v[1];
    float mean = 0.f, invstd = 1.f, g = 1.f, be = 0.f;
    if (gamma != nullptr) {
        const double s0 = warp_sum_d((has0 ? (double)v[0] : 0.0) + (has1 ? (double)v[1] : 0.0));
        const double s1 = warp_sum_d((has0 ? (double)v[0] * v[0] : 0.0) + (has1 ? (double)v[1] * v[1] : 0.0));
        const double mu = s0 / M;
        double var = s1 / M - mu * mu;
        if (var < 0.0) var = 0.0;
        mean = (float)mu; invstd = (float)(1.0 / sqrt(var + (double)eps));
        g = gamma[c]; be = beta[c];
        if (lane == 0) {
            mean_o[c] = mean; invstd_o[c] = invstd;
            if (running_mean) {
                const double unbiased = stat_count > 1.0 ? var * stat_count / (stat_count - 1.0) : var;
                running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
                running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
            }
        }
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int m = lane + 32 * h;
        if (m < M) {
            const float o = (v[h] - mean) * invstd * g + be;
            y[(size_t)m * C + c] = relu ? fmaxf(o, 0.f) : o;
        }
    }
}

// Backward of the layer for one channel per warp: dy -> (ReLU mask from y) -> BatchNorm backward over the M rows -> dy_raw
// (rows lane, lane + 32 in registers; written to dy_raw_out [M][C] for the input-gradient kernel), parameter gradients
// db / dgamma / dbeta accumulated, then dw[c,l] += sum_m dy_raw[m] z[m,l] with the lanes across l (coalesced z rows).
__global__ void __launch_bounds__(256)
vec_layer_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ z, const float* __restrict__ y_raw,
                     const float* __restrict__ y, int M, int L, int C, const float* __restrict__ gamma,
                     const float* __restrict__ mean, const float* __restrict__ invstd, int relu, float* __restrict__ dw,
                     float* __restrict__ db, float* __restrict__ dgamma, float* __restrict__ dbeta,
                     float* __restrict__ dy_raw_out) {
    const int lane = threadIdx.x & 31, c = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (c >= C) return;                                   // warp-uniform
    float d[2] = {0.f, 0.f};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int m = lane + 32 * h;
        if (m < M) {
            float g = dy[(size_t)m * C + c];
            if (relu && !(y[(size_t)m * C + c] > 0.f)) g = 0.f;
            d[h] = g;
        }
    }
    if (gamma != nullptr) {
        const float mu = mean[c], is = invstd[c];
        float xh[2] = {0.f, 0.f};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int m = lane + 32 * h;
            if (m < M) xh[h] = (y_raw[(size_t)m * C + c] - mu) * is;
        }
        const double s0 = warp_sum_d((double)d[0] + (double)d[1]);
        const double s1 = warp_sum_d((double)d[0] * xh[0] + (double)d[1] * xh[1]);
        if (lane == 0 && dgamma) { dgamma[c] += (float)s1; dbeta[c] += (float)s0; }
        const float m0 = (float)(s0 / M), m1 = (float)(s1 / M), a = gamma[c] * is;
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (lane + 32 * h < M) d[h] = a * (d[h] - m0 - xh[h] * m1);
    }
    if (db != nullptr) {
        const float s = warp_sum(d[0] + d[1]);
        if (lane == 0) db[c] += s;
    }
    if (dy_raw_out != nullptr) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (lane + 32 * h < M) dy_raw_out[(size_t)(lane + 32 * h) * C + c] = d[h];
    }
    if (dw != nullptr) {
        for (int l0 = 0; l0 < L; l0 += 128) {             // 4 columns per lane per pass (unconditional, clamped loads)
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            int li[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) li[u] = min(l0 + lane + 32 * u, L - 1);
#pragma unroll 4
            for (int m = 0; m < M; ++m) {
                const float g = __shfl_sync(0xffffffffu, m < 32 ? d[0] : d[1], m & 31);
                const float* zr = z + (size_t)m * L;
#pragma unroll
                for (int u = 0; u < 4; ++u) acc[u] = fmaf(g, __ldg(zr + li[u]), acc[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (l0 + lane + 32 * u < L) dw[(size_t)c * L + l0 + lane + 32 * u] += acc[u];
        }
    }
}

// dz[m,l] += sum_c dy_raw[m,c] w[c,l] from the [M][C] pre-BN gradient the kernel above wrote. One block per 32 columns l
// (lane = column); the 8 warps take the channels c = warp, warp + 8, ... and their partial sums are combined through shared
// memory in warp order - a fixed order without atomics: the input gradient is bit-reproducible from run to run (it feeds the
// BatchNorm backward of the 2x2 / 4x4 encoder maps, which amplifies fp32 atomics-order noise by orders of magnitude).
__global__ void __launch_bounds__(256)
vec_layer_dz_kernel(const float* __restrict__ dy_raw, const float* __restrict__ w, int M, int L, int C,
                    float* __restrict__ dz) {
    __shared__ float part[8][16][33];
    const int lane = threadIdx.x & 31, wg = threadIdx.x >> 5;
    const int l = blockIdx.x * 32 + lane;
    for (int m0 = 0; m0 < M; m0 += 16) {                  // 16 rows per pass
        float acc[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) acc[j] = 0.f;
        const int lc = min(l, L - 1);
        int row[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) row[j] = min(m0 + j, M - 1);          // clamped: loads stay unconditional
#pragma unroll 2
        for (int c = wg; c < C; c += 8) {
            const float wv = __ldg(w + (size_t)c * L + lc);
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[j] = fmaf(__ldg(dy_raw + (size_t)row[j] * C + c), wv, acc[j]);   // warp-uniform address
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) part[wg][j][lane] = acc[j];
        __syncthreads();
        for (int i = threadIdx.x; i < 16 * 32; i += 256) {
            const int j = i >> 5, ll = i & 31;
            if (m0 + j < M && blockIdx.x * 32 + ll < L) {
                float t = 0.f;
#pragma unroll
                for (int g = 0; g < 8; ++g) t += part[g][j][ll];
                dz[(size_t)(m0 + j) * L + blockIdx.x * 32 + ll] += t;
            }
        }
        __syncthreads();
    }
}

// z = mu + eps * exp(.5 lv) * T  ->  dmu += dz (summed over samples), dlv += dz * eps * .5 * exp(.5 lv) * T
__global__ void reparam_bwd_kernel(const float* __restrict__ dz, const float* __restrict__ eps, const float* __restrict__ lv,
                                   float T, float* __restrict__ dmu, float* __restrict__ dlv, size_t n, int ns) {
    for (size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x; j < n; j += (size_t)gridDim.x * blockDim.x) {
        float a = 0.f, b = 0.f;
        const float h = 0.5f * expf(0.5f * lv[j]) * T;
        for (int s = 0; s < ns; ++s) { const float g = dz[s * n + j]; a += g; b = fmaf(g, eps[s * n + j] * h, b); }
        dmu[j] += a; dlv[j] += b;
    }
}

// ---------------------------------------------------------------- stem 7x7 s2 p3 weight gradient (Cin <= 4)
// dw[co][ci][r][s] = sum_{n,ho,wo} dy[n,ho,wo,co] * x[n,ci,2ho+r-3,2wo+s-3]. Block = one 8x8 output tile of one image:
// the input patch (Cin x 21 x 21) and the dy tile (64 px x Cout) sit in smem, thread t owns filter entries t, t+256, ...
__global__ void __launch_bounds__(256)
stem_wgrad_kernel(const float* __restrict__ x, const uint16_t* __restrict__ dy, int dy_ctot, int dy_coff,
                  float* __restrict__ dw, int Cin, int H, int W, int Cout, int Ho, int Wo, int dtype) {
    constexpr int TO = 8, K = 7, TI = (TO - 1) * 2 + K;   // 21
    extern __shared__ float sm[];
    float* s_in = sm;                           // [Cin][TI][TI]
    float* s_dy = sm + 4 * TI * TI;             // [64 px][Cout]
    const int n = blockIdx.z, oh0 = blockIdx.y * TO, ow0 = blockIdx.x * TO;
    for (int i = threadIdx.x; i < Cin * TI * TI; i += 256) {
        const int c = i / (TI * TI), r = (i / TI) % TI, q = i % TI;
        const int ih = oh0 * 2 - 3 + r, iw = ow0 * 2 - 3 + q;
        s_in[i] = (ih >= 0 && ih < H && iw >= 0 && iw < W) ? x[(((size_t)n * Cin + c) * H + ih) * W + iw] : 0.f;
    }
    for (int i = threadIdx.x; i < TO * TO * Cout; i += 256) {
        const int px = i / Cout, co = i % Cout;
        const int oh = oh0 + px / TO, ow = ow0 + px % TO;
        s_dy[i] = (oh < Ho && ow < Wo) ? h2f(dy[(((size_t)n * Ho + oh) * Wo + ow) * dy_ctot + dy_coff + co], dtype) : 0.f;
    }
    __syncthreads();
    const int total = Cout * Cin * K * K;
    for (int e = threadIdx.x; e < total; e += 256) {
        const int co = e / (Cin * K * K), rem = e % (Cin * K * K);
        const int ci = rem / (K * K), r = (rem / K) % K, s = rem % K;
        float acc = 0.f;
#pragma unroll 8
        for (int px = 0; px < TO * TO; ++px)
            acc = fmaf(s_dy[px * Cout + co], s_in[(ci * TI + (px / TO) * 2 + r) * TI + (px % TO) * 2 + s], acc);
        atomicAdd(&dw[e], acc);
    }
}

// ---------------------------------------------------------------- gradient norm + AdamW
__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ g, size_t n, double* __restrict__ out) {
    float a = 0.f;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        a = fmaf(g[i], g[i], a);
    a = warp_sum(a);
    __shared__ float sh[8];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int w = 0; w < 8; ++w) t += sh[w];
        atomicAdd(out, (double)t);
    }
}

// torch.optim.AdamW (decoupled weight decay) after clip_grad_norm_(max_norm): the clip coefficient is computed on the
// device from the squared norm so the step needs no host synchronisation.
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, size_t n,
             float lr, float beta1, float beta2, float eps, float wd, float bc1, float bc2, const double* __restrict__ sumsq,
             float max_norm, float grad_scale) {
    float clip = 1.0f;
    if (sumsq != nullptr && max_norm > 0.f) {
        const float norm = sqrtf((float)sumsq[0]) * grad_scale;
        clip = fminf(1.0f, max_norm / (norm + 1e-6f));
    }
    const float gs = clip * grad_scale;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float gi = g[i] * gs;
        float pi = p[i] * (1.f - lr * wd);
        const float mi = beta1 * m[i] + (1.f - beta1) * gi;
        const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;
        m[i] = mi; v[i] = vi;
        pi -= lr * (mi / bc1) / (sqrtf(vi / bc2) + eps);
        p[i] = pi;
    }
}

static int ts_grid(size_t n) {
    size_t g = (n + 255) / 256;
    const size_t cap = (size_t)num_sms() * 8;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

}  // namespace vu

using namespace vu;

extern "C" int vu_map_bn_sigmoid_fwd(const float* x, size_t n, const float* gamma, const float* beta, float eps, float momentum,
                                     float* running_mean, float* running_var, double* sums, float* stat, float* psi,
                                     vu_stream_t st) {
    if (!x || !sums || !stat || !psi || !gamma || !beta || n < 1) return set_error(VU_ERR_INVALID, "map_bn_sigmoid_fwd: bad args");
    cudaStream_t s = (cudaStream_t)st;
    cudaMemsetAsync(sums, 0, 2 * sizeof(double), s);
    map_sums_kernel<0><<<ts_grid(n), 256, 0, s>>>(x, nullptr, nullptr, nullptr, n, sums);
    map_bn_finalize_kernel<<<1, 1, 0, s>>>(sums, (double)n, eps, momentum, stat, running_mean, running_var);
    map_bn_sigmoid_apply_kernel<<<ts_grid(n), 256, 0, s>>>(x, stat, gamma, beta, psi, n);
    VU_CHECK_LAUNCH("map_bn_sigmoid_fwd");
    return 0;
}

extern "C" int vu_map_bn_sigmoid_bwd(const float* x, const float* dpsi, const float* psi, const float* stat, const float* gamma,
                                     size_t n, double* sums, float* dx, float* dgamma, float* dbeta, vu_stream_t st) {
    if (!x || !dpsi || !psi || !stat || !gamma || !sums || !dx || !dgamma || !dbeta) return set_error(VU_ERR_INVALID, "map_bn_sigmoid_bwd: null");
    cudaStream_t s = (cudaStream_t)st;
    cudaMemsetAsync(sums, 0, 2 * sizeof(double), s);
    map_sums_kernel<1><<<ts_grid(n), 256, 0, s>>>(x, dpsi, psi, stat, n, sums);
    map_bn_sigmoid_bwd_kernel<<<ts_grid(n), 256, 0, s>>>(x, dpsi, psi, stat, gamma, sums, (double)n, dx, dgamma, dbeta, n);
    VU_CHECK_LAUNCH("map_bn_sigmoid_bwd");
    return 0;
}

extern "C" int vu_map_sum(const float* x, size_t n, double* sums, vu_stream_t st) {
    if (!x || !sums) return set_error(VU_ERR_INVALID, "map_sum: null");
    cudaStream_t s = (cudaStream_t)st;
    cudaMemsetAsync(sums, 0, 2 * sizeof(double), s);
    map_sums_kernel<0><<<ts_grid(n), 256, 0, s>>>(x, nullptr, nullptr, nullptr, n, sums);
    VU_CHECK_LAUNCH("map_sum");
    return 0;
}

extern "C" int vu_vec_layer_fwd(const float* z, const float* w, const float* b, int M, int L, int C, const float* gamma,
                                const float* beta, float eps, float momentum, double stat_count, float* running_mean,
                                float* running_var, int relu, float* y_raw, float* y, float* mean, float* invstd,
                                vu_stream_t st) {
    if (!z || !w || !y_raw || !y || M < 1 || M > kVecM) return set_error(VU_ERR_INVALID, "vec_layer_fwd: bad args (M <= 64)");
    if (gamma && (!beta || !mean || !invstd)) return set_error(VU_ERR_INVALID, "vec_layer_fwd: BN needs beta/mean/invstd");
    vec_layer_fwd_kernel<<<(C + 7) / 8, 256, 0, (cudaStream_t)st>>>(z, w, b, M, L, C, gamma, beta, eps, momentum, stat_count,
                                                                      running_mean, running_var, relu, y_raw, y, mean, invstd);
    VU_CHECK_LAUNCH("vec_layer_fwd");
    return 0;
}

extern "C" int vu_vec_layer_bwd(const float* dy, const float* z, const float* w, const float* y_raw, const float* y, int M,
                                int L, int C, const float* gamma, const float* mean, const float* invstd, int relu, float* dz,
                                float* dw, float* db, float* dgamma, float* dbeta, float* scratch, vu_stream_t st) {
    if (!dy || !z || !w || !y_raw || !y || M < 1 || M > kVecM) return set_error(VU_ERR_INVALID, "vec_layer_bwd: bad args (M <= 64)");
    if (dz && !scratch) return set_error(VU_ERR_INVALID, "vec_layer_bwd: dz needs the M*C scratch buffer");
    vec_layer_bwd_kernel<<<(C + 7) / 8, 256, 0, (cudaStream_t)st>>>(dy, z, y_raw, y, M, L, C, gamma, mean, invstd, relu, dw,
                                                                      db, dgamma, dbeta, dz ? scratch : nullptr);
    VU_CHECK_LAUNCH("vec_layer_bwd");
    if (dz != nullptr) {
        vec_layer_dz_kernel<<<(L + 31) / 32, 256, 0, (cudaStream_t)st>>>(scratch, w, M, L, C, dz);
        VU_CHECK_LAUNCH("vec_layer_dz");
    }
    return 0;
}

extern "C" int vu_reparam_bwd(const float* dz, const float* eps, const float* logvar, float temperature, float* dmu,
                              float* dlogvar, size_t n_per_sample, int num_samples, vu_stream_t st) {
    if (!dz || !eps || !logvar || !dmu || !dlogvar) return set_error(VU_ERR_INVALID, "reparam_bwd: null");
    reparam_bwd_kernel<<<ts_grid(n_per_sample), 256, 0, (cudaStream_t)st>>>(dz, eps, logvar, temperature, dmu, dlogvar,
                                                                          n_per_sample, num_samples);
    VU_CHECK_LAUNCH("reparam_bwd");
    return 0;
}

extern "C" int vu_stem_wgrad(const float* x_nchw, const void* dy, int dy_ctot, int dy_coff, float* dw, int N, int Cin, int H,
                             int W, int Cout, int dtype, vu_stream_t st) {
    if (!x_nchw || !dy || !dw || Cin < 1 || Cin > 4 || Cout > 64) return set_error(VU_ERR_INVALID, "stem_wgrad: bad args (Cin <= 4, Cout <= 64)");
    const int Ho = (H + 6 - 7) / 2 + 1, Wo = (W + 6 - 7) / 2 + 1;
    const size_t smem = (4 * 21 * 21 + 64 * Cout) * sizeof(float);
    dim3 grid((Wo + 7) / 8, (Ho + 7) / 8, N);
    stem_wgrad_kernel<<<grid, 256, smem, (cudaStream_t)st>>>(x_nchw, (const uint16_t*)dy, dy_ctot, dy_coff, dw, Cin, H, W, Cout,
                                                            Ho, Wo, dtype);
    VU_CHECK_LAUNCH("stem_wgrad");
    return 0;
}

extern "C" int vu_sumsq(const float* g, size_t n, double* out, int zero_first, vu_stream_t st) {
    if (!g || !out) return set_error(VU_ERR_INVALID, "sumsq: null");
    cudaStream_t s = (cudaStream_t)st;
    if (zero_first) cudaMemsetAsync(out, 0, sizeof(double), s);
    sumsq_kernel<<<ts_grid(n), 256, 0, s>>>(g, n, out);
    VU_CHECK_LAUNCH("sumsq");
    return 0;
}

extern "C" int vu_adamw_step(float* p, const float* g, float* m, float* v, size_t n, float lr, float beta1, float beta2,
                             float eps, float weight_decay, int step, const double* grad_sumsq, float max_norm,
                             float grad_scale, vu_stream_t st) {
    if (!p || !g || !m || !v || step < 1) return set_error(VU_ERR_INVALID, "adamw_step: bad args");
    const float bc1 = 1.0f - powf(beta1, (float)step), bc2 = 1.0f - powf(beta2, (float)step);
    adamw_kernel<<<ts_grid(n), 256, 0, (cudaStream_t)st>>>(p, g, m, v, n, lr, beta1, beta2, eps, weight_decay, bc1, bc2,
                                                           grad_sumsq, max_norm, grad_scale);
    VU_CHECK_LAUNCH("adamw_step");
    return 0;
}
