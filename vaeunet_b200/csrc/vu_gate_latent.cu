// Attention-gate tail, latent head (GAP + mu/logvar + reparameterisation), constant-map
// projections and the 1x1 output head. All memory/latency bound: 16-byte accesses, shuffles.
#include "vu_internal.h"
#include "vu_half.cuh"

namespace vu {

// G threads cooperate on one pixel (G = power of two <= 32, G*8 >= max(F, Cl) unless looping).
template <int G>
__global__ void __launch_bounds__(256)
attention_gate_kernel(const uint16_t* __restrict__ g1, int g1_ctot, int g1_coff, const uint16_t* __restrict__ x1,
                      int x1_ctot, int x1_coff, const uint16_t* __restrict__ skip, int skip_ctot, int skip_coff,
                      const float* __restrict__ psi_w, float psi_scale, float psi_bias, uint16_t* __restrict__ out,
                      int out_ctot, int out_coff, float* __restrict__ psi_out, size_t npix, int rep, int HW, int F8,
                      int C8, int dtype) {
    const int sub = threadIdx.x % G;
    const size_t gid = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) / G;
    const size_t gstride = ((size_t)gridDim.x * blockDim.x) / G;
    const size_t gl = (threadIdx.x & 31) / G;  // group index inside the warp (loop exit is warp-uniform)
    for (size_t pix_ = gid; pix_ - gl < npix; pix_ += gstride) {
        const bool valid = pix_ < npix;
        const size_t pix = valid ? pix_ : npix - 1;
        const size_t n = pix / HW, hw = pix % HW;
        const size_t spix = (n / rep) * HW + hw;
        float dot = 0.0f;
        for (int f8 = sub; f8 < F8; f8 += G) {
            const uint4 a = __ldg(reinterpret_cast<const uint4*>(g1 + pix * g1_ctot + g1_coff) + f8);
            const uint4 b = __ldg(reinterpret_cast<const uint4*>(x1 + spix * x1_ctot + x1_coff) + f8);
            float fa[8], fb[8];
            unpack8(a, fa, dtype); unpack8(b, fb, dtype);
            const float4 w0 = __ldg(reinterpret_cast<const float4*>(psi_w) + f8 * 2);
            const float4 w1 = __ldg(reinterpret_cast<const float4*>(psi_w) + f8 * 2 + 1);
            const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
            for (int j = 0; j < 8; ++j) dot = fmaf(w[j], fmaxf(fa[j] + fb[j], 0.0f), dot);
        }
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o, G);
        const float psi = 1.0f / (1.0f + expf(-(psi_scale * dot + psi_bias)));
        if (!valid) continue;
        if (psi_out != nullptr && sub == 0) psi_out[pix] = psi;
        for (int c8 = sub; c8 < C8; c8 += G) {
            const uint4 s = __ldg(reinterpret_cast<const uint4*>(skip + spix * skip_ctot + skip_coff) + c8);
            float f[8];
            unpack8(s, f, dtype);
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] *= psi;
            *(reinterpret_cast<uint4*>(out + pix * out_ctot + out_coff) + c8) = pack8(f, dtype);
        }
    }
}

// one CTA per image: GAP over h*w into smem, then 2L dot products (warp per output), then z for all samples
__global__ void __launch_bounds__(256)
latent_head_kernel(const uint16_t* __restrict__ f4, int HW, int C, int f_ctot, int f_coff,
                   const float* __restrict__ w_mu, const float* __restrict__ b_mu, const float* __restrict__ w_lv,
                   const float* __restrict__ b_lv, const float* __restrict__ eps, float temperature, int Ns, int B,
                   int L, float* __restrict__ mu, float* __restrict__ logvar, float* __restrict__ z, int dtype) {
    extern __shared__ float sm[];
    float* s_mean = sm;          // [C]
    float* s_out = sm + C;       // [2L]
    float* s_part = sm + C + 2 * L;   // [lanes_per][C] per-lane partial sums
    const int b = blockIdx.x;
    const int C8 = C / 8;
    const int lanes_per = blockDim.x / C8 > 0 ? blockDim.x / C8 : 1;
    // phase 1: thread t owns channel group t % C8 and strides over pixels; the lanes' partials are combined in a FIXED order
    // (smem atomics made mu / logvar differ by an ulp from run to run, which every decode then amplifies through its
    // 16-bit roundings: sharded and unsharded MC could not agree bit for bit)
    {
        const int c8 = threadIdx.x % C8;
        const int pl = threadIdx.x / C8;
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (pl < lanes_per) {
            for (int p = pl; p < HW; p += lanes_per) {
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(f4 + ((size_t)b * HW + p) * f_ctot + f_coff) + c8);
                float f[8];
                unpack8(v, f, dtype);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] += f[j];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) s_part[pl * C + c8 * 8 + j] = acc[j];
        }
    }
    __syncthreads();
    const float inv = 1.0f / (float)HW;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float t = 0.0f;
        for (int l = 0; l < lanes_per; ++l) t += s_part[l * C + c];
        s_mean[c] = t * inv;
    }
    __syncthreads();
    // phase 2
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (int o = warp; o < 2 * L; o += nwarps) {
        const float* wrow = (o < L) ? (w_mu + (size_t)o * C) : (w_lv + (size_t)(o - L) * C);
        float d = 0.0f;
        for (int c = lane; c < C; c += 32) d = fmaf(__ldg(wrow + c), s_mean[c], d);
        d = warp_sum(d);
        if (lane == 0) s_out[o] = d + ((o < L) ? b_mu[o] : b_lv[o - L]);
    }
    __syncthreads();
    for (int l = threadIdx.x; l < L; l += blockDim.x) {
        mu[(size_t)b * L + l] = s_out[l];
        logvar[(size_t)b * L + l] = s_out[L + l];
    }
    if (z != nullptr) {
        for (int i = threadIdx.x; i < Ns * L; i += blockDim.x) {
            const int n = i / L, l = i % L;
            const float m = s_out[l], lv = s_out[L + l];
            float v = m;
            if (eps != nullptr) {
                // utils/vae_utils.py:8-10: std = exp(.5 lv) * T ; z = mu + eps * std
                const float sd = expf(0.5f * lv) * temperature;
                v = m + eps[((size_t)n * B + b) * L + l] * sd;
            }
            z[((size_t)n * B + b) * L + l] = v;
        }
    }
}

__global__ void latent_project_kernel(const float* __restrict__ z, const float* __restrict__ w,
                                      const float* __restrict__ scale, const float* __restrict__ shift,
                                      float* __restrict__ v, int M, int L, int C, int relu) {
    const size_t total = (size_t)M * C;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int c = idx % C;
        const int i = idx / C;
        float d = 0.0f;
        for (int l = 0; l < L; ++l) d = fmaf(__ldg(w + (size_t)c * L + l), __ldg(z + (size_t)i * L + l), d);
        d = d * (scale ? scale[c] : 1.0f) + (shift ? shift[c] : 0.0f);
        v[idx] = relu ? fmaxf(d, 0.0f) : d;
    }
}

// logits[pix] = <w, x[pix, :]> + bias ; G = C/8 threads per pixel
template <int G>
__global__ void __launch_bounds__(256)
final_conv1_kernel(const uint16_t* __restrict__ x, size_t npix, int x_ctot, int x_coff, const float* __restrict__ w,
                   float bias, const float* __restrict__ bias_dev, float* __restrict__ logits, int C8, int dtype) {
    if (bias_dev) bias += bias_dev[0];
    const int sub = threadIdx.x % G;
    const size_t gid = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) / G;
    const size_t gstride = ((size_t)gridDim.x * blockDim.x) / G;
    const size_t gl = (threadIdx.x & 31) / G;
    for (size_t pix_ = gid; pix_ - gl < npix; pix_ += gstride) {
        const bool valid = pix_ < npix;
        const size_t pix = valid ? pix_ : npix - 1;
        float d = 0.0f;
        for (int c8 = sub; c8 < C8; c8 += G) {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(x + pix * x_ctot + x_coff) + c8);
            float f[8];
            unpack8(v, f, dtype);
            const float4 w0 = __ldg(reinterpret_cast<const float4*>(w) + c8 * 2);
            const float4 w1 = __ldg(reinterpret_cast<const float4*>(w) + c8 * 2 + 1);
            d = fmaf(f[0], w0.x, d); d = fmaf(f[1], w0.y, d); d = fmaf(f[2], w0.z, d); d = fmaf(f[3], w0.w, d);
            d = fmaf(f[4], w1.x, d); d = fmaf(f[5], w1.y, d); d = fmaf(f[6], w1.z, d); d = fmaf(f[7], w1.w, d);
        }
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o, G);
        if (sub == 0 && valid) logits[pix] = d + bias;
    }
}

// pos_bias table: out[d][cls][co] = sum_{(r,s) inside the image for class cls} sum_c w[co][coff+c][r][s] * z[d][c]
// (row class 0 = first output row: tap row r = 0 falls in the zero padding; 2 = last row: r = 2 does; same for columns).
__global__ void __launch_bounds__(256)
border_bias_kernel(const float* __restrict__ w, int Cout, int Cin_total, int coff, int Cz, const float* __restrict__ z, int D,
                   float* __restrict__ out) {
    const size_t total = (size_t)D * 9 * Cout;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int co = idx % Cout, cls = (idx / Cout) % 9, d = idx / ((size_t)Cout * 9);
        const int rc = cls / 3, cc = cls % 3;
        const int r_lo = rc == 0 ? 1 : 0, r_hi = rc == 2 ? 1 : 2, s_lo = cc == 0 ? 1 : 0, s_hi = cc == 2 ? 1 : 2;
        float acc = 0.f;
        for (int c = 0; c < Cz; ++c) {
            const float* wp = w + ((size_t)co * Cin_total + coff + c) * 9;
            float t = 0.f;
            for (int r = r_lo; r <= r_hi; ++r)
                for (int q = s_lo; q <= s_hi; ++q) t += wp[r * 3 + q];
            acc = fmaf(t, z[(size_t)d * Cz + c], acc);
        }
        out[idx] = acc;
    }
}

// The same in two steps for weights that stay fixed (inference): wsum[cls][co][c] = sum of the in-image taps of
// w[co][coff+c], built once; then out[d][cls][co] = <wsum[cls][co][:], z[d][:]> per call (Cz % 4 == 0, 16-byte rows).
__global__ void __launch_bounds__(256)
border_wsum_kernel(const float* __restrict__ w, int Cout, int Cin_total, int coff, int Cz, float* __restrict__ wsum) {
    const size_t total = (size_t)9 * Cout * Cz;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int c = idx % Cz, co = (idx / Cz) % Cout, cls = idx / ((size_t)Cz * Cout);
        const int rc = cls / 3, cc = cls % 3;
        const int r_lo = rc == 0 ? 1 : 0, r_hi = rc == 2 ? 1 : 2, s_lo = cc == 0 ? 1 : 0, s_hi = cc == 2 ? 1 : 2;
        const float* wp = w + ((size_t)co * Cin_total + coff + c) * 9;
        float t = 0.f;
        for (int r = r_lo; r <= r_hi; ++r)
            for (int q = s_lo; q <= s_hi; ++q) t += wp[r * 3 + q];
        wsum[idx] = t;
    }
}

__global__ void __launch_bounds__(256)
border_bias_apply_kernel(const float* __restrict__ wsum, const float* __restrict__ z, int D, int Cout, int Cz,
                         float* __restrict__ out) {
    const size_t total = (size_t)D * 9 * Cout;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int co = idx % Cout, cls = (idx / Cout) % 9, d = idx / ((size_t)Cout * 9);
        const float4* wp = reinterpret_cast<const float4*>(wsum + ((size_t)cls * Cout + co) * Cz);
        const float4* zp = reinterpret_cast<const float4*>(z + (size_t)d * Cz);
        float acc = 0.f;
        for (int c = 0; c < Cz / 4; ++c) {
            const float4 a = __ldg(wp + c), b = __ldg(zp + c);
            acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc); acc = fmaf(a.z, b.z, acc); acc = fmaf(a.w, b.w, acc);
        }
        out[idx] = acc;
    }
}

// Tiled form for wide constant slices (block 0 of the decoder: the 512 channels of z_initial(z)): per border class a
// [Cout x Cz] x [Cz x D] product, 64 x 32 outputs per block through shared memory, 4 x 2 per thread.
__global__ void __launch_bounds__(256)
border_bias_gemm_kernel(const float* __restrict__ wsum, const float* __restrict__ z, int D, int Cout, int Cz,
                        float* __restrict__ out, int accumulate) {
    __shared__ float As[64][33];
    __shared__ float Zs[32][33];
    const int cls = blockIdx.z, co0 = blockIdx.x * 64, d0 = blockIdx.y * 32;
    const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
    float acc[4][2] = {};
    for (int c0 = 0; c0 < Cz; c0 += 32) {
        for (int i = threadIdx.x; i < 64 * 32; i += 256) {
            const int r = i / 32, c = i % 32;
            As[r][c] = (co0 + r < Cout && c0 + c < Cz) ? wsum[((size_t)cls * Cout + co0 + r) * Cz + c0 + c] : 0.f;
        }
        for (int i = threadIdx.x; i < 32 * 32; i += 256) {
            const int r = i / 32, c = i % 32;
            Zs[r][c] = (d0 + r < D && c0 + c < Cz) ? z[(size_t)(d0 + r) * Cz + c0 + c] : 0.f;
        }
        __syncthreads();
#pragma unroll 8
        for (int k = 0; k < 32; ++k) {
            float a[4], b[2];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[tx * 4 + i][k];
#pragma unroll
            for (int j = 0; j < 2; ++j) b[j] = Zs[ty * 2 + j][k];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 2; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const int d = d0 + ty * 2 + j;
        if (d >= D) continue;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int co = co0 + tx * 4 + i;
            if (co >= Cout) continue;
            float* o = out + ((size_t)d * 9 + cls) * Cout + co;
            *o = accumulate ? *o + acc[i][j] : acc[i][j];
        }
    }
}

__global__ void reparameterize_kernel(const float* __restrict__ mu, const float* __restrict__ lv,
                                      const float* __restrict__ eps, float temperature, float* __restrict__ z,
                                      size_t n_per_sample, size_t total) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t j = i % n_per_sample;
        z[i] = mu[j] + eps[i] * (expf(0.5f * lv[j]) * temperature);
    }
}

static int grid_cap(size_t blocks) {
    const size_t cap = (size_t)num_sms() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace vu

using namespace vu;

extern "C" int vu_attention_gate(const void* g1, int g1_ctot, int g1_coff, const void* x1, int x1_ctot, int x1_coff,
                                 const void* skip, int skip_ctot, int skip_coff, const float* psi_w, float psi_scale,
                                 float psi_bias, void* out, int out_ctot, int out_coff, float* psi, int Ng, int rep,
                                 int H, int W, int F, int Cl, int dtype, vu_stream_t st) {
    if (!g1 || !x1 || !skip || !psi_w || !out) return set_error(VU_ERR_INVALID, "attention_gate: null");
    if (rep < 1 || Ng % rep) return set_error(VU_ERR_INVALID, "attention_gate: Ng (%d) must be a multiple of rep (%d)", Ng, rep);
    if (F % 8 || Cl % 8 || g1_ctot % 8 || g1_coff % 8 || x1_ctot % 8 || x1_coff % 8 || skip_ctot % 8 || skip_coff % 8 ||
        out_ctot % 8 || out_coff % 8)
        return set_error(VU_ERR_ALIGN, "attention_gate: channels %% 8");
    const size_t npix = (size_t)Ng * H * W;
    const int mx = (F > Cl ? F : Cl) / 8;
    cudaStream_t s = (cudaStream_t)st;
#define VU_GATE(G)                                                                                                   \
    attention_gate_kernel<G><<<grid_cap((npix * G + 255) / 256), 256, 0, s>>>(                                        \
        (const uint16_t*)g1, g1_ctot, g1_coff, (const uint16_t*)x1, x1_ctot, x1_coff, (const uint16_t*)skip, skip_ctot, \
        skip_coff, psi_w, psi_scale, psi_bias, (uint16_t*)out, out_ctot, out_coff, psi, npix, rep, H* W, F / 8, Cl / 8, dtype)
    if (mx <= 4) VU_GATE(4);
    else if (mx <= 8) VU_GATE(8);
    else if (mx <= 16) VU_GATE(16);
    else VU_GATE(32);
#undef VU_GATE
    VU_CHECK_LAUNCH("attention_gate");
    return 0;
}

extern "C" int vu_latent_head(const void* f4, int B, int h, int w, int C, int f_ctot, int f_coff, const float* w_mu,
                              const float* b_mu, const float* w_lv, const float* b_lv, const float* eps,
                              float temperature, int Ns, int L, float* mu, float* logvar, float* z, int dtype,
                              vu_stream_t st) {
    if (!f4 || !w_mu || !b_mu || !w_lv || !b_lv || !mu || !logvar) return set_error(VU_ERR_INVALID, "latent_head: null");
    if (C % 8 || f_ctot % 8 || f_coff % 8) return set_error(VU_ERR_ALIGN, "latent_head: channels %% 8");
    if (C / 8 > 256) return set_error(VU_ERR_UNSUPPORTED, "latent_head: C <= 2048");
    if (B < 1 || L < 1 || Ns < 0) return set_error(VU_ERR_INVALID, "latent_head: bad sizes");
    const int lanes_per = 256 / (C / 8) > 0 ? 256 / (C / 8) : 1;
    const size_t smem = (size_t)(C + 2 * L + (size_t)lanes_per * C) * sizeof(float);
    if (smem > 48 * 1024) return set_error(VU_ERR_UNSUPPORTED, "latent_head: C / L too large for shared memory");
    latent_head_kernel<<<B, 256, smem, (cudaStream_t)st>>>((const uint16_t*)f4, h * w, C, f_ctot, f_coff, w_mu, b_mu, w_lv,
                                                           b_lv, eps, temperature, Ns, B, L, mu, logvar, z, dtype);
    VU_CHECK_LAUNCH("latent_head");
    return 0;
}

extern "C" int vu_latent_project(const float* z, const float* w, const float* scale, const float* shift, float* v, int M,
                                 int L, int C, int relu, vu_stream_t st) {
    if (!z || !w || !v || M < 1) return set_error(VU_ERR_INVALID, "latent_project: bad args");
    const size_t total = (size_t)M * C;
    latent_project_kernel<<<grid_cap((total + 255) / 256), 256, 0, (cudaStream_t)st>>>(z, w, scale, shift, v, M, L, C, relu);
    VU_CHECK_LAUNCH("latent_project");
    return 0;
}

extern "C" int vu_final_conv1(const void* x, int N, int H, int W, int C, int x_ctot, int x_coff, const float* w,
                              float bias, const float* bias_dev, float* logits, int dtype, vu_stream_t st) {
    if (!x || !w || !logits) return set_error(VU_ERR_INVALID, "final_conv1: null");
    if (C % 8 || x_ctot % 8 || x_coff % 8) return set_error(VU_ERR_ALIGN, "final_conv1: channels %% 8");
    const size_t npix = (size_t)N * H * W;
    const int c8 = C / 8;
    cudaStream_t s = (cudaStream_t)st;
    if (c8 <= 4)
        final_conv1_kernel<4><<<grid_cap((npix * 4 + 255) / 256), 256, 0, s>>>((const uint16_t*)x, npix, x_ctot, x_coff, w, bias, bias_dev, logits, c8, dtype);
    else if (c8 <= 8)
        final_conv1_kernel<8><<<grid_cap((npix * 8 + 255) / 256), 256, 0, s>>>((const uint16_t*)x, npix, x_ctot, x_coff, w, bias, bias_dev, logits, c8, dtype);
    else
        final_conv1_kernel<16><<<grid_cap((npix * 16 + 255) / 256), 256, 0, s>>>((const uint16_t*)x, npix, x_ctot, x_coff, w, bias, bias_dev, logits, c8, dtype);
    VU_CHECK_LAUNCH("final_conv1");
    return 0;
}

extern "C" int vu_border_bias(const float* w_oihw, int Cout, int Cin_total, int coff, int Cz, const float* z, int D, float* out,
                              vu_stream_t st) {
    if (!w_oihw || !z || !out || Cout < 1 || Cz < 1 || D < 1 || coff < 0 || coff + Cz > Cin_total)
        return set_error(VU_ERR_INVALID, "border_bias: bad args");
    const size_t total = (size_t)D * 9 * Cout;
    border_bias_kernel<<<grid_cap((total + 255) / 256), 256, 0, (cudaStream_t)st>>>(w_oihw, Cout, Cin_total, coff, Cz, z, D, out);
    VU_CHECK_LAUNCH("border_bias");
    return 0;
}

extern "C" int vu_border_wsum(const float* w_oihw, int Cout, int Cin_total, int coff, int Cz, float* wsum, vu_stream_t st) {
    if (!w_oihw || !wsum || Cout < 1 || Cz < 1 || coff < 0 || coff + Cz > Cin_total)
        return set_error(VU_ERR_INVALID, "border_wsum: bad args");
    const size_t total = (size_t)9 * Cout * Cz;
    border_wsum_kernel<<<grid_cap((total + 255) / 256), 256, 0, (cudaStream_t)st>>>(w_oihw, Cout, Cin_total, coff, Cz, wsum);
    VU_CHECK_LAUNCH("border_wsum");
    return 0;
}

extern "C" int vu_border_bias_apply(const float* wsum, const float* z, int D, int Cout, int Cz, float* out, int accumulate,
                                    vu_stream_t st) {
    if (!wsum || !z || !out || D < 1 || Cout < 1 || Cz < 4 || Cz % 4) return set_error(VU_ERR_INVALID, "border_bias_apply: bad args (Cz % 4)");
    const size_t total = (size_t)D * 9 * Cout;
    if (Cz >= 128 || accumulate) {
        dim3 grid((Cout + 63) / 64, (D + 31) / 32, 9);
        border_bias_gemm_kernel<<<grid, 256, 0, (cudaStream_t)st>>>(wsum, z, D, Cout, Cz, out, accumulate);
    } else
        border_bias_apply_kernel<<<grid_cap((total + 255) / 256), 256, 0, (cudaStream_t)st>>>(wsum, z, D, Cout, Cz, out);
    VU_CHECK_LAUNCH("border_bias_apply");
    return 0;
}

extern "C" int vu_reparameterize(const float* mu, const float* logvar, const float* eps, float temperature, float* z,
                                 size_t n_per_sample, int num_samples, vu_stream_t st) {
    if (!mu || !logvar || !eps || !z || num_samples < 1) return set_error(VU_ERR_INVALID, "reparameterize: bad args");
    const size_t total = n_per_sample * (size_t)num_samples;
    reparameterize_kernel<<<grid_cap((total + 255) / 256), 256, 0, (cudaStream_t)st>>>(mu, logvar, eps, temperature, z,
                                                                                      n_per_sample, total);
    VU_CHECK_LAUNCH("reparameterize");
    return 0;
}
