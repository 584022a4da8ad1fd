// Segmentation losses, KL with free bits, calibration / confusion reductions.
// One pass over the maps, warp-shuffle + one fp64 atomic per CTA per accumulator.
#include "vu_internal.h"
#include "vu_half.cuh"

namespace vu {

__device__ __forceinline__ float nan_to_num_(float v, float nanv, float posinf, float neginf) {
    if (isnan(v)) return nanv;
    if (isinf(v)) return v > 0 ? posinf : neginf;
    return v;
}

template <int NACC>
__device__ __forceinline__ void block_accumulate(double (&v)[NACC], double* __restrict__ out) {
    __shared__ double s_part[8][NACC];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < NACC; ++i) v[i] = warp_sum_d(v[i]);
    if (lane == 0)
#pragma unroll
        for (int i = 0; i < NACC; ++i) s_part[warp][i] = v[i];
    __syncthreads();
    if (threadIdx.x < NACC) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_part[w][threadIdx.x];
        atomicAdd(out + threadIdx.x, t);
    }
}

__global__ void __launch_bounds__(256)
seg_loss_sums_kernel(const float* __restrict__ x, const float* __restrict__ t, size_t n, int is_prob, float alpha,
                     float gamma, double* __restrict__ sums) {
    double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float xv = x[i], tv = t[i];
        float p, bce;
        if (is_prob) {
            p = xv;
            bce = 0.0f;
        } else {
            p = 1.0f / (1.0f + expf(-xv));
            // F.binary_cross_entropy_with_logits: max(x,0) - x*t + log1p(exp(-|x|))
            bce = fmaxf(xv, 0.0f) - xv * tv + log1pf(expf(-fabsf(xv)));
        }
        const float ps = nan_to_num_(p, 0.0f, 1.0f, 0.0f);           // utils/loss.py:12-14,79
        const float pt = tv * ps + (1.0f - tv) * (1.0f - ps);
        const float fw = powf(1.0f - pt, gamma);
        const float at = tv * alpha + (1.0f - tv) * (1.0f - alpha);
        const float bce_e = -tv * logf(ps + 1e-6f) - (1.0f - tv) * logf(1.0f - ps + 1e-6f);
        const float focal = nan_to_num_(at * fw * bce_e, 0.0f, 1.0f, 0.0f);
        const float pc = fminf(fmaxf(p, 1e-9f), 1.0f - 1e-9f);        // uncertainty_metrics.py:336
        const float nll = -(tv * logf(pc) + (1.0f - tv) * logf(1.0f - pc));
        acc[0] += ps; acc[1] += tv; acc[2] += ps * tv; acc[3] += bce; acc[4] += focal; acc[5] += 1.0;
        acc[6] += (double)(p - tv) * (p - tv); acc[7] += nll;
    }
    block_accumulate<8>(acc, sums);
}

__global__ void __launch_bounds__(256)
seg_loss_bwd_kernel(const float* __restrict__ x, const float* __restrict__ t, size_t n, const double* __restrict__ sums,
                    float w_bce, float w_dice, float w_focal, float alpha, float gamma, float smooth, float gscale,
                    float* __restrict__ dx) {
    const float P = (float)sums[0], T = (float)sums[1], I = (float)sums[2];
    const float half = smooth * 0.5f;
    const float Pc = fmaxf(P, half), Tc = fmaxf(T, half);
    const float D = Pc + Tc + smooth;
    const float num = 2.0f * I + smooth;
    const float dP = (P >= half) ? 1.0f : 0.0f;  // clamp(min) passes the gradient when input >= min
    const float inv_n = 1.0f / (float)n;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float xv = x[i], tv = t[i];
        const float p = 1.0f / (1.0f + expf(-xv));
        const float dpdx = p * (1.0f - p);
        float g = 0.0f;
        if (w_bce != 0.0f) g += w_bce * (p - tv) * inv_n;
        if (w_dice != 0.0f) {
            // dice = 1 - num/D ; d/dp_i = -(2 t_i D - num dP) / D^2
            g += w_dice * (-(2.0f * tv * D - num * dP) / (D * D)) * dpdx;
        }
        if (w_focal != 0.0f) {
            const float pt = tv * p + (1.0f - tv) * (1.0f - p);
            const float at = tv * alpha + (1.0f - tv) * (1.0f - alpha);
            const float om = 1.0f - pt;
            const float bce_e = -tv * logf(p + 1e-6f) - (1.0f - tv) * logf(1.0f - p + 1e-6f);
            const float dbce = -tv / (p + 1e-6f) + (1.0f - tv) / (1.0f - p + 1e-6f);
            const float dfw = (om > 0.0f) ? -gamma * powf(om, gamma - 1.0f) * (2.0f * tv - 1.0f) : 0.0f;
            const float df = at * (dfw * bce_e + powf(om, gamma) * dbce);
            g += w_focal * df * dpdx * inv_n;
        }
        dx[i] = g * gscale;
    }
}

// one CTA; B*L is tiny (16 x 32)
__global__ void kl_free_bits_kernel(const float* __restrict__ mu, const float* __restrict__ lv, int B, int L, float fb,
                                    float* __restrict__ kl_out, float* __restrict__ dmu, float* __restrict__ dlv,
                                    float gscale) {
    __shared__ float s_red[32];
    float acc = 0.0f;
    for (int i = threadIdx.x; i < B * L; i += blockDim.x) {
        float m = mu[i], l = lv[i];
        if (isnan(m)) m = 0.0f;                      // nan_to_num(nan=0.0), utils/loss.py:152-153
        if (isnan(l)) l = 0.0f;
        m = nan_to_num_(m, 0.0f, 3.4028234663852886e38f, -3.4028234663852886e38f);
        l = nan_to_num_(l, 0.0f, 3.4028234663852886e38f, -3.4028234663852886e38f);
        const float e = expf(l);
        float k = 0.5f * (m * m + e - l - 1.0f);
        const bool in_clamp = (k >= -100.0f) && (k <= 100.0f);
        k = fminf(fmaxf(k, -100.0f), 100.0f);
        bool live = in_clamp;
        if (fb > 0.0f) { if (k < fb) { k = fb; live = false; } }
        acc += k;
        if (dmu) dmu[i] = live ? gscale * m / (float)B : 0.0f;
        if (dlv) dlv[i] = live ? gscale * 0.5f * (e - 1.0f) / (float)B : 0.0f;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_red[w];
        t /= (float)B;
        if (isnan(t)) t = 1e-8f;                     // utils/loss.py:168
        *kl_out = t;
    }
}

struct BinBounds { float b[66]; };

__global__ void __launch_bounds__(256)
ece_hist_kernel(const float* __restrict__ p, const float* __restrict__ gt, size_t n, int nb, BinBounds bb,
                double* __restrict__ hist) {
    extern __shared__ float s_h[];  // [3*nb] : count, sum gt, sum p  (fp32 per CTA, fp64 globally)
    for (int i = threadIdx.x; i < 3 * nb; i += blockDim.x) s_h[i] = 0.0f;
    __syncthreads();
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float pv = p[i];
        int bin = -1;
        for (int b = 0; b < nb; ++b)
            if (pv >= bb.b[b] && pv < bb.b[b + 1]) { bin = b; break; }
        if (bin >= 0) {
            atomicAdd(&s_h[bin], 1.0f);
            atomicAdd(&s_h[nb + bin], gt[i]);
            atomicAdd(&s_h[2 * nb + bin], pv);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 3 * nb; i += blockDim.x) atomicAdd(hist + i, (double)s_h[i]);
}

__global__ void __launch_bounds__(256)
confusion_kernel(const float* __restrict__ x, const float* __restrict__ t, size_t n, float thr, double* __restrict__ counts) {
    double acc[4] = {0, 0, 0, 0};
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const bool pp = x[i] > thr, tt = t[i] > 0.5f;
        acc[0] += (pp && tt); acc[1] += (pp && !tt); acc[2] += (!pp && !tt); acc[3] += (!pp && tt);
    }
    block_accumulate<4>(acc, counts);
}

static int grid_for_loss(size_t n) {
    size_t g = (n + 255) / 256 / 8 + 1;
    const size_t cap = (size_t)num_sms() * 8;
    if (g > cap) g = cap;
    return (int)g;
}

}  // namespace vu

using namespace vu;

extern "C" int vu_seg_loss_sums(const float* x, const float* t, size_t n, int is_prob, float alpha, float gamma,
                                double* sums, vu_stream_t st) {
    if (!x || !t || !sums || n == 0) return set_error(VU_ERR_INVALID, "seg_loss_sums: bad args");
    cudaStream_t s = (cudaStream_t)st;
    cudaError_t e = cudaMemsetAsync(sums, 0, 8 * sizeof(double), s);
    if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "memset: %s", cudaGetErrorString(e));
    seg_loss_sums_kernel<<<grid_for_loss(n), 256, 0, s>>>(x, t, n, is_prob, alpha, gamma, sums);
    VU_CHECK_LAUNCH("seg_loss_sums");
    return 0;
}

extern "C" int vu_seg_loss_bwd(const float* x, const float* t, size_t n, const double* sums, float w_bce, float w_dice,
                               float w_focal, float alpha, float gamma, float smooth, float gscale, float* dx,
                               vu_stream_t st) {
    if (!x || !t || !sums || !dx || n == 0) return set_error(VU_ERR_INVALID, "seg_loss_bwd: bad args");
    seg_loss_bwd_kernel<<<grid_for_loss(n) * 4, 256, 0, (cudaStream_t)st>>>(x, t, n, sums, w_bce, w_dice, w_focal, alpha,
                                                                            gamma, smooth, gscale, dx);
    VU_CHECK_LAUNCH("seg_loss_bwd");
    return 0;
}

extern "C" int vu_kl_free_bits(const float* mu, const float* lv, int B, int L, float fb, float* kl_out, float* dmu,
                               float* dlv, float gscale, vu_stream_t st) {
    if (!mu || !lv || !kl_out || B < 1 || L < 1) return set_error(VU_ERR_INVALID, "kl_free_bits: bad args");
    kl_free_bits_kernel<<<1, 256, 0, (cudaStream_t)st>>>(mu, lv, B, L, fb, kl_out, dmu, dlv, gscale);
    VU_CHECK_LAUNCH("kl_free_bits");
    return 0;
}

extern "C" int vu_ece_hist(const float* p, const float* gt, size_t n, int nb, double* hist, vu_stream_t st) {
    if (!p || !gt || !hist || nb < 1 || nb > 64) return set_error(VU_ERR_INVALID, "ece_hist: bad args (1..64 bins)");
    cudaStream_t s = (cudaStream_t)st;
    BinBounds bb{};
    // torch.linspace(0, 1, nb+1) in fp32 (ATen fills symmetrically from both ends)
    const int steps = nb + 1;
    const float step = 1.0f / (float)(steps - 1);
    for (int k = 0; k < steps; ++k) bb.b[k] = (k < steps / 2) ? step * k : 1.0f - step * (float)(steps - 1 - k);
    cudaError_t e = cudaMemsetAsync(hist, 0, 3 * nb * sizeof(double), s);
    if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "memset: %s", cudaGetErrorString(e));
    ece_hist_kernel<<<grid_for_loss(n), 256, 3 * nb * sizeof(float), s>>>(p, gt, n, nb, bb, hist);
    VU_CHECK_LAUNCH("ece_hist");
    return 0;
}

extern "C" int vu_confusion_counts(const float* x, const float* t, size_t n, float thr, double* counts, vu_stream_t st) {
    if (!x || !t || !counts || n == 0) return set_error(VU_ERR_INVALID, "confusion_counts: bad args");
    cudaStream_t s = (cudaStream_t)st;
    cudaError_t e = cudaMemsetAsync(counts, 0, 4 * sizeof(double), s);
    if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "memset: %s", cudaGetErrorString(e));
    confusion_kernel<<<grid_for_loss(n), 256, 0, s>>>(x, t, n, thr, counts);
    VU_CHECK_LAUNCH("confusion_counts");
    return 0;
}
