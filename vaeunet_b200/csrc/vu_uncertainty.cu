// Final upsample (+sigmoid), per-pixel MC uncertainty reduction (Welford), Chan merge of
// sample shards, and ramp-weighted tile blending. All fp32, HBM-bound streaming kernels.
#include "vu_internal.h"
#include "vu_half.cuh"

namespace vu {

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

struct Lerp { int i0, i1; float l0, l1; };
__device__ __forceinline__ Lerp make_lerp(float scale, int dst, int in_size) {
    const float r = scale * dst;
    Lerp L;
    L.i0 = (int)r;
    L.i1 = L.i0 + ((L.i0 < in_size - 1) ? 1 : 0);
    L.l1 = r - L.i0;
    L.l0 = 1.0f - L.l1;
    return L;
}
__device__ __forceinline__ float bilerp(const float* __restrict__ m, int Ws, const Lerp& y, const Lerp& x, bool pre_sig) {
    float a = __ldg(m + (size_t)y.i0 * Ws + x.i0), b = __ldg(m + (size_t)y.i0 * Ws + x.i1);
    float c = __ldg(m + (size_t)y.i1 * Ws + x.i0), d = __ldg(m + (size_t)y.i1 * Ws + x.i1);
    if (pre_sig) { a = sigmoidf_(a); b = sigmoidf_(b); c = sigmoidf_(c); d = sigmoidf_(d); }
    return y.l0 * (x.l0 * a + x.l1 * b) + y.l1 * (x.l0 * c + x.l1 * d);
}

__global__ void resize_logits_kernel(const float* __restrict__ x, float* __restrict__ y, int N, int Hs, int Ws, int Ho,
                                     int Wo, float sh, float sw, int mode) {
    const size_t total = (size_t)N * Ho * Wo;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int wo = idx % Wo;
        const int ho = (idx / Wo) % Ho;
        const size_t n = idx / ((size_t)Wo * Ho);
        float v;
        if (Hs == Ho && Ws == Wo) {
            v = x[idx];
            if (mode != 0) v = sigmoidf_(v);
        } else {
            const Lerp ly = make_lerp(sh, ho, Hs), lx = make_lerp(sw, wo, Ws);
            v = bilerp(x + n * Hs * Ws, Ws, ly, lx, mode == 1);
            if (mode == 2) v = sigmoidf_(v);
        }
        y[idx] = v;
    }
}

__device__ __forceinline__ float bin_entropy(float p) {
    const float e = 1e-7f;
    return -(p * logf(p + e) + (1.0f - p) * logf(1.0f - p + e));
}

// logits [B][Ns][Hs][Ws] -> per-pixel statistics over Ns (visualize_vae.py:90-117).
__global__ void __launch_bounds__(256)
uncertainty_reduce_kernel(const float* __restrict__ in, int B, int Ns, int Hs, int Ws, int Ho, int Wo, int is_prob,
                          float sh, float sw, float* __restrict__ mean_o, float* __restrict__ std_o,
                          float* __restrict__ ent_o, float* __restrict__ mi_o, float* __restrict__ cv_o,
                          float* __restrict__ probs_o) {
    const size_t total = (size_t)B * Ho * Wo;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int wo = idx % Wo;
        const int ho = (idx / Wo) % Ho;
        const size_t b = idx / ((size_t)Wo * Ho);
        Lerp ly, lx;
        if (!is_prob) { ly = make_lerp(sh, ho, Hs); lx = make_lerp(sw, wo, Ws); }
        float mean = 0.0f, m2 = 0.0f, sum_h = 0.0f;
        const bool need_h = mi_o != nullptr;       // per-sample entropies only feed the mutual information
        for (int n = 0; n < Ns; ++n) {
            float p;
            if (is_prob) {
                p = __ldg(in + ((b * Ns + n) * Ho + ho) * Wo + wo);
            } else {
                const float* m = in + (b * Ns + n) * (size_t)Hs * Ws;
                const float lg = (Hs == Ho && Ws == Wo) ? __ldg(m + (size_t)ho * Ws + wo) : bilerp(m, Ws, ly, lx, false);
                p = sigmoidf_(lg);
                if (probs_o) probs_o[((b * Ns + n) * Ho + ho) * Wo + wo] = p;
            }
            const float d = p - mean;             // Welford
            mean += d * __frcp_rn((float)(n + 1));
            m2 += d * (p - mean);
            if (need_h) sum_h += bin_entropy(p);
        }
        const float var = (Ns > 1) ? m2 / (float)(Ns - 1) : __int_as_float(0x7fc00000);  // torch.std: NaN for N=1
        const float sd = sqrtf(fmaxf(var, 0.0f));
        const float sd_out = (Ns > 1) ? sd : var;
        const float ent = bin_entropy(mean);
        if (mean_o) mean_o[idx] = mean;
        if (std_o) std_o[idx] = sd_out;
        if (ent_o) ent_o[idx] = ent;
        if (mi_o) mi_o[idx] = ent - sum_h / (float)Ns;
        if (cv_o) cv_o[idx] = sd_out / (mean + 1e-7f);
    }
}

__global__ void moments_partial_kernel(const float* __restrict__ probs, int Ns, size_t npix, float* __restrict__ mean_o,
                                       float* __restrict__ m2_o, float* __restrict__ sh_o) {
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < npix; idx += (size_t)gridDim.x * blockDim.x) {
        float mean = 0.0f, m2 = 0.0f, sum_h = 0.0f;
        for (int n = 0; n < Ns; ++n) {
            const float p = __ldg(probs + (size_t)n * npix + idx);
            const float d = p - mean;
            mean += d * __frcp_rn((float)(n + 1));
            m2 += d * (p - mean);
            sum_h += bin_entropy(p);
        }
        mean_o[idx] = mean; m2_o[idx] = m2; sh_o[idx] = sum_h;
    }
}

struct ShardCounts { float c[64]; };

__global__ void moments_finalize_kernel(ShardCounts cnt, const float* __restrict__ mean_s, const float* __restrict__ m2_s,
                                        const float* __restrict__ sh_s, int nshards, size_t npix,
                                        float* __restrict__ mean_o, float* __restrict__ std_o, float* __restrict__ ent_o,
                                        float* __restrict__ mi_o, float* __restrict__ cv_o) {
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < npix; idx += (size_t)gridDim.x * blockDim.x) {
        float n = 0.0f, mean = 0.0f, m2 = 0.0f, sh = 0.0f;
        for (int k = 0; k < nshards; ++k) {
            const float nb = cnt.c[k];
            if (nb <= 0.0f) continue;
            const float mb = mean_s[(size_t)k * npix + idx], m2b = m2_s[(size_t)k * npix + idx];
            const float nn = n + nb, d = mb - mean;   // Chan et al. pairwise merge
            mean += d * (nb / nn);
            m2 += m2b + d * d * (n * nb / nn);
            n = nn;
            sh += sh_s[(size_t)k * npix + idx];
        }
        const float var = (n > 1.0f) ? m2 / (n - 1.0f) : __int_as_float(0x7fc00000);
        const float sd = (n > 1.0f) ? sqrtf(fmaxf(var, 0.0f)) : var;
        const float ent = bin_entropy(mean);
        if (mean_o) mean_o[idx] = mean;
        if (std_o) std_o[idx] = sd;
        if (ent_o) ent_o[idx] = ent;
        if (mi_o) mi_o[idx] = ent - sh / n;
        if (cv_o) cv_o[idx] = sd / (mean + 1e-7f);
    }
}

// torch.linspace(0, 1, n)[k] in fp32 (ATen: symmetric fill from both ends)
__device__ __forceinline__ float linspace01(int k, int n) {
    if (n <= 1) return 0.0f;
    const float step = 1.0f / (float)(n - 1);
    return (k < n / 2) ? step * k : 1.0f - step * (float)(n - 1 - k);
}

__global__ void blend_accumulate_kernel(const float* __restrict__ tile, int ph, int pw, int sh, int sw, int ti, int tj,
                                        int nth, int ntw, int ov, float* __restrict__ acc, float* __restrict__ wsum, int W) {
    const size_t total = (size_t)ph * pw;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int x = idx % pw, y = idx / pw;
        float w = 1.0f;
        if (ph > 2 * ov) {  // visualize_vae.py:364-378 (axis 2 first, then axis 3)
            if (ti > 0 && y < ov) w *= linspace01(y, ov);
            if (ti < nth - 1 && y >= ph - ov) w *= (1.0f - linspace01(y - (ph - ov), ov));
        }
        if (pw > 2 * ov) {
            if (tj > 0 && x < ov) w *= linspace01(x, ov);
            if (tj < ntw - 1 && x >= pw - ov) w *= (1.0f - linspace01(x - (pw - ov), ov));
        }
        const size_t o = (size_t)(sh + y) * W + sw + x;
        acc[o] += tile[idx] * w;
        wsum[o] += w;
    }
}

__global__ void blend_finalize_kernel(const float* __restrict__ acc, const float* __restrict__ wsum, float* __restrict__ out,
                                      size_t n) {
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x)
        out[idx] = acc[idx] / (wsum[idx] + 1e-8f);
}

static int grid_for_n(size_t total) {
    size_t g = (total + 255) / 256;
    const size_t cap = (size_t)num_sms() * 16;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

}  // namespace vu

using namespace vu;

extern "C" int vu_resize_logits(const float* x, float* y, int N, int Hs, int Ws, int Ho, int Wo, int mode, vu_stream_t st) {
    if (!x || !y || N < 1) return set_error(VU_ERR_INVALID, "resize_logits: bad args");
    const float sh = Ho > 1 ? (float)(Hs - 1) / (float)(Ho - 1) : 0.0f;
    const float sw = Wo > 1 ? (float)(Ws - 1) / (float)(Wo - 1) : 0.0f;
    resize_logits_kernel<<<grid_for_n((size_t)N * Ho * Wo), 256, 0, (cudaStream_t)st>>>(x, y, N, Hs, Ws, Ho, Wo, sh, sw, mode);
    VU_CHECK_LAUNCH("resize_logits");
    return 0;
}

extern "C" int vu_uncertainty_reduce(const float* logits, int B, int Ns, int Hs, int Ws, int Ho, int Wo, int is_prob,
                                     float* mean, float* stdev, float* entropy, float* mi, float* cv, float* probs_out,
                                     vu_stream_t st) {
    if (!logits || B < 1 || Ns < 1) return set_error(VU_ERR_INVALID, "uncertainty_reduce: bad args");
    if (is_prob && (Hs != Ho || Ws != Wo)) return set_error(VU_ERR_INVALID, "uncertainty_reduce: probability input must be at output size");
    const float sh = Ho > 1 ? (float)(Hs - 1) / (float)(Ho - 1) : 0.0f;
    const float sw = Wo > 1 ? (float)(Ws - 1) / (float)(Wo - 1) : 0.0f;
    uncertainty_reduce_kernel<<<grid_for_n((size_t)B * Ho * Wo), 256, 0, (cudaStream_t)st>>>(
        logits, B, Ns, Hs, Ws, Ho, Wo, is_prob, sh, sw, mean, stdev, entropy, mi, cv, probs_out);
    VU_CHECK_LAUNCH("uncertainty_reduce");
    return 0;
}

extern "C" int vu_moments_partial(const float* probs, int Ns, size_t npix, float* mean, float* m2, float* sum_ent,
                                  vu_stream_t st) {
    if (!probs || !mean || !m2 || !sum_ent || Ns < 0) return set_error(VU_ERR_INVALID, "moments_partial: bad args");
    moments_partial_kernel<<<grid_for_n(npix), 256, 0, (cudaStream_t)st>>>(probs, Ns, npix, mean, m2, sum_ent);
    VU_CHECK_LAUNCH("moments_partial");
    return 0;
}

extern "C" int vu_moments_finalize(const float* counts_host, const float* mean, const float* m2, const float* sum_ent,
                                   int nshards, size_t npix, float* out_mean, float* out_std, float* out_entropy,
                                   float* out_mi, float* out_cv, vu_stream_t st) {
    if (!counts_host || !mean || !m2 || !sum_ent || nshards < 1 || nshards > 64)
        return set_error(VU_ERR_INVALID, "moments_finalize: bad args (1..64 shards)");
    ShardCounts c{};
    for (int k = 0; k < nshards; ++k) c.c[k] = counts_host[k];
    moments_finalize_kernel<<<grid_for_n(npix), 256, 0, (cudaStream_t)st>>>(c, mean, m2, sum_ent, nshards, npix, out_mean,
                                                                           out_std, out_entropy, out_mi, out_cv);
    VU_CHECK_LAUNCH("moments_finalize");
    return 0;
}

extern "C" int vu_blend_accumulate(const float* tile, int ph, int pw, int sh, int sw, int ti, int tj, int nth, int ntw,
                                   int overlap, float* acc, float* wsum, int H, int W, vu_stream_t st) {
    if (!tile || !acc || !wsum) return set_error(VU_ERR_INVALID, "blend_accumulate: null");
    if (sh < 0 || sw < 0 || sh + ph > H || sw + pw > W) return set_error(VU_ERR_INVALID, "blend_accumulate: tile outside image");
    if (overlap < 0) return set_error(VU_ERR_INVALID, "blend_accumulate: negative overlap");
    blend_accumulate_kernel<<<grid_for_n((size_t)ph * pw), 256, 0, (cudaStream_t)st>>>(tile, ph, pw, sh, sw, ti, tj, nth, ntw,
                                                                                      overlap, acc, wsum, W);
    VU_CHECK_LAUNCH("blend_accumulate");
    return 0;
}

extern "C" int vu_blend_finalize(const float* acc, const float* wsum, float* out, size_t n, vu_stream_t st) {
    if (!acc || !wsum || !out) return set_error(VU_ERR_INVALID, "blend_finalize: null");
    blend_finalize_kernel<<<grid_for_n(n), 256, 0, (cudaStream_t)st>>>(acc, wsum, out, n);
    VU_CHECK_LAUNCH("blend_finalize");
    return 0;
}
