// Fused producer of a DecoderBlock's concat buffer (unet/unet_resnet.py:71-98):
//
//   cat[d, ho, wo, :] = [ bilinear_ac(x)[d,ho,wo,:]  |  skip * sigmoid(BN(psi . relu(g1 + x1)))  |  z_proj(z)[d,:] ]
//
// in ONE pass over the output (bound: HBM writes of `cat`). What it replaces in the unfused path:
// bilinear_ac_nhwc_kernel + the W_g 1x1 conv at OUTPUT resolution + attention_gate_kernel +
// broadcast_vec_nhwc_kernel, each a full read/write of a high-resolution map per MC sample.
//
// W_g is a 1x1 conv (+ folded BN: an affine map per pixel) and bilinear interpolation weights sum
// to 1, so  BN(W_g(upsample(x))) == upsample(BN(W_g(x)))  exactly in real arithmetic: g1 is computed
// by the tensor-core conv at the LOW resolution of x (4x fewer pixels) and interpolated here on the
// fly. A spatially constant x (bottleneck z_initial map) is a 1x1 source map.
//
// Mapping: 16 lanes per output pixel (2 pixels per warp), every lane moves 16-byte (8-channel) chunks.
#include <cstdlib>
#include "vu_half.cuh"
#include "vu_internal.h"

namespace vu {

struct CatParams {
    const uint16_t* x; int hs, ws, cx8, x_ctot, x_coff;          // low-resolution decoder input [D,hs,ws,Cx]
    const uint16_t* g1; int g1_ctot, g1_coff;                     // BN(W_g x) at low resolution [D,hs,ws,F]
    const uint16_t* x1; int x1_ctot, x1_coff;                     // BN(W_x skip) [Dp,H,W,F], Dp = D / rep
    const uint16_t* skip; int skip_ctot, skip_coff;               // [Dp,H,W,Cs]
    int f8, cs8;
    const float* psi_w; float psi_scale, psi_bias;
    const float* zvec; int cz8;                                   // [D, Cz] fp32 or null
    uint16_t* out; int out_ctot;
    float* psi_out;
    int D, rep, H, W;
    float sh, sw;
    int dtype;
};

__device__ __forceinline__ uint4 lerp4(const uint4& a, const uint4& b, const uint4& c, const uint4& d, float h0l,
                                       float h1l, float w0l, float w1l, int dtype, float (&o)[8]) {
    float fa[8], fb[8], fc[8], fd[8];
    unpack8(a, fa, dtype); unpack8(b, fb, dtype); unpack8(c, fc, dtype); unpack8(d, fd, dtype);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = h0l * (w0l * fa[j] + w1l * fb[j]) + h1l * (w0l * fc[j] + w1l * fd[j]);   // ATen's order
    return pack8(o, dtype);
}

// One warp = 8 consecutive output pixels of one row; every phase (gate, x slice, skip slice, z slice) runs
// over the flattened (pixel, 16-byte chunk) list of those 8 pixels so all 32 lanes stay busy and several
// independent loads are in flight per lane. Per-pixel interpolation coordinates and psi go through smem.
constexpr int kCatPix = 8;          // pixels per warp
constexpr int kCatWarps = 8;        // warps per block -> 64 consecutive pixels of a row per block iteration
constexpr int kCatMaxAcc = 4;       // gate accumulators per lane: F/8 * 8 pixels / 32 lanes = F/32 <= 4

__device__ __forceinline__ void split_idx(int idx, int n, int shift, int& px, int& c8) {
    if (shift >= 0) { px = idx >> shift; c8 = idx & (n - 1); }
    else { px = idx / n; c8 = idx - px * n; }
}

// 16-byte chunk helpers with the storage type as a template parameter (no per-element runtime switch).
// fp16 storage interpolates / scales in packed half2 arithmetic (4 HFMA2 per 4 values instead of convert +
// fp32 FMA + convert: the kernel is issue-bound otherwise); bf16 storage keeps fp32 arithmetic.
template <int DT>
__device__ __forceinline__ uint4 lerp_chunk(const uint4& a, const uint4& b, const uint4& c, const uint4& d, float h0l,
                                            float h1l, float w0l, float w1l) {
    if constexpr (DT == 1) {
        const __half2 w00 = __float2half2_rn(h0l * w0l), w01 = __float2half2_rn(h0l * w1l);
        const __half2 w10 = __float2half2_rn(h1l * w0l), w11 = __float2half2_rn(h1l * w1l);
        const __half2* pa = reinterpret_cast<const __half2*>(&a);
        const __half2* pb = reinterpret_cast<const __half2*>(&b);
        const __half2* pc = reinterpret_cast<const __half2*>(&c);
        const __half2* pd = reinterpret_cast<const __half2*>(&d);
        uint4 r;
        __half2* pr = reinterpret_cast<__half2*>(&r);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            pr[j] = __hfma2(w11, pd[j], __hfma2(w10, pc[j], __hfma2(w01, pb[j], __hmul2(w00, pa[j]))));
        return r;
    } else {
        float o[8];
        return lerp4(a, b, c, d, h0l, h1l, w0l, w1l, 0, o);
    }
}
template <int DT>
__device__ __forceinline__ uint4 scale_chunk(const uint4& v, float s) {
    if constexpr (DT == 1) {
        const __half2 s2 = __float2half2_rn(s);
        const __half2* pv = reinterpret_cast<const __half2*>(&v);
        uint4 r;
        __half2* pr = reinterpret_cast<__half2*>(&r);
#pragma unroll
        for (int j = 0; j < 4; ++j) pr[j] = __hmul2(pv[j], s2);
        return r;
    } else {
        float f[8];
        unpack8(v, f, 0);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] *= s;
        return pack8(f, 0);
    }
}

template <int DT>
__global__ void __launch_bounds__(kCatWarps * 32)
decoder_concat_kernel(const CatParams p, int segs_per_row, uint32_t num_segs, int sh_f, int sh_x, int sh_s, int sh_z) {
    __shared__ int s_w1[kCatWarps][kCatPix];
    __shared__ int s_w1p[kCatWarps][kCatPix];
    __shared__ float s_w1l[kCatWarps][kCatPix];
    __shared__ float s_psi[kCatWarps][kCatPix];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t HW = (uint32_t)p.H * p.W;
    for (uint32_t seg = blockIdx.x; seg < num_segs; seg += gridDim.x) {
        // seg enumerates (patch, ho, 64-pixel segment, sample): the `rep` decodes of one patch take the same skip / x1 rows,
        // so they are adjacent in the grid-stride order and those rows are read from DRAM once, not once per decode
        // (ncu at 160 decodes: 1.61 GB read for 0.70 GB of distinct inputs with the decode-major order)
        const uint32_t r = seg % (uint32_t)p.rep;
        const uint32_t q = seg / (uint32_t)p.rep;
        const uint32_t prow = q / segs_per_row;                  // = patch * H + ho
        const int wseg = q - prow * segs_per_row;
        const uint32_t pb = prow / p.H;
        const int ho = prow - pb * p.H;
        const uint32_t d = pb * p.rep + r;
        const uint32_t row = d * p.H + ho;
        const int wo0 = (wseg * kCatWarps + warp) * kCatPix;
        if (wo0 >= p.W) continue;                                // warp-uniform
        const int npx = min(kCatPix, p.W - wo0);
        // ATen upsample_bilinear2d, align_corners=True: src = dst * (in-1)/(out-1) in fp32
        const float h1r = p.sh * ho;
        const int h1 = (int)h1r;
        const int h1p = (h1 < p.hs - 1) ? 1 : 0;
        const float h1l = h1r - h1, h0l = 1.0f - h1l;
        __syncwarp();
        if (lane < kCatPix) {
            const float w1r = p.sw * (wo0 + lane);
            const int w1 = min((int)w1r, p.ws - 1);
            s_w1[warp][lane] = w1;
            s_w1p[warp][lane] = (w1 < p.ws - 1) ? 1 : 0;
            s_w1l[warp][lane] = w1r - (int)w1r;
        }
        __syncwarp();
        const size_t srow0 = ((size_t)d * p.hs + h1) * p.ws;     // first source pixel of the upper source row
        const size_t srow1 = srow0 + (size_t)h1p * p.ws;
        const size_t spix0 = (size_t)(d / p.rep) * HW + (size_t)ho * p.W + wo0;   // skip / x1 pixel of px = 0
        const size_t opix0 = (size_t)row * p.W + wo0;

        // ---- phase 1: psi[px] = sigmoid(scale * <psi_w, relu(upsample(g1) + x1)> + bias)
        {
            float acc[kCatMaxAcc] = {0.f, 0.f, 0.f, 0.f};
            const uint16_t* gb = p.g1 + p.g1_coff;
#pragma unroll
            for (int k = 0; k < kCatMaxAcc; ++k) {
                const int idx = lane + 32 * k;
                if (idx < kCatPix * p.f8) {
                    int px, f8;
                    split_idx(idx, p.f8, sh_f, px, f8);
                    if (px < npx) {
                        const int w1 = s_w1[warp][px], w1p = s_w1p[warp][px];
                        const float w1l = s_w1l[warp][px], w0l = 1.0f - w1l;
                        float g[8], xb[8];
                        lerp4(__ldg(reinterpret_cast<const uint4*>(gb + (srow0 + w1) * p.g1_ctot) + f8),
                              __ldg(reinterpret_cast<const uint4*>(gb + (srow0 + w1 + w1p) * p.g1_ctot) + f8),
                              __ldg(reinterpret_cast<const uint4*>(gb + (srow1 + w1) * p.g1_ctot) + f8),
                              __ldg(reinterpret_cast<const uint4*>(gb + (srow1 + w1 + w1p) * p.g1_ctot) + f8), h0l, h1l, w0l,
                              w1l, DT, g);
                        unpack8(__ldg(reinterpret_cast<const uint4*>(p.x1 + (spix0 + px) * p.x1_ctot + p.x1_coff) + f8), xb, DT);
                        const float4 wa = __ldg(reinterpret_cast<const float4*>(p.psi_w) + f8 * 2);
                        const float4 wb = __ldg(reinterpret_cast<const float4*>(p.psi_w) + f8 * 2 + 1);
                        const float w[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
                        float dsum = 0.0f;
#pragma unroll
                        for (int j = 0; j < 8; ++j) dsum = fmaf(w[j], fmaxf(g[j] + xb[j], 0.0f), dsum);
                        acc[k] = dsum;
                    }
                }
            }
            // lanes that share a pixel: f8 <= 32 -> f8 consecutive lanes (one accumulator per 32/f8 pixels);
            // f8 > 32 is rejected on the host
            const int grp = min(p.f8, 32);
#pragma unroll
            for (int k = 0; k < kCatMaxAcc; ++k) {
                float v = acc[k];
                for (int o = grp >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                const int idx = lane + 32 * k;
                if (idx < kCatPix * p.f8 && (lane & (grp - 1)) == 0) {
                    const int px = idx / p.f8;
                    const float psi = 1.0f / (1.0f + expf(-(p.psi_scale * v + p.psi_bias)));
                    s_psi[warp][px] = psi;
                    if (p.psi_out != nullptr && px < npx) p.psi_out[opix0 + px] = psi;
                }
            }
        }
        __syncwarp();

        // ---- phase 2: slice 0 = upsampled x
        {
            const uint16_t* xb = p.x + p.x_coff;
            const int total = kCatPix * p.cx8;
            for (int idx = lane; idx < total; idx += 32) {
                int px, c8;
                split_idx(idx, p.cx8, sh_x, px, c8);
                if (px >= npx) continue;
                const int w1 = s_w1[warp][px], w1p = s_w1p[warp][px];
                const float w1l = s_w1l[warp][px], w0l = 1.0f - w1l;
                const uint4 v = lerp_chunk<DT>(__ldg(reinterpret_cast<const uint4*>(xb + (srow0 + w1) * p.x_ctot) + c8),
                                               __ldg(reinterpret_cast<const uint4*>(xb + (srow0 + w1 + w1p) * p.x_ctot) + c8),
                                               __ldg(reinterpret_cast<const uint4*>(xb + (srow1 + w1) * p.x_ctot) + c8),
                                               __ldg(reinterpret_cast<const uint4*>(xb + (srow1 + w1 + w1p) * p.x_ctot) + c8),
                                               h0l, h1l, w0l, w1l);
                *(reinterpret_cast<uint4*>(p.out + (opix0 + px) * p.out_ctot) + c8) = v;
            }
        }
        // ---- phase 3: slice 1 = gated skip
        {
            const int total = kCatPix * p.cs8;
            for (int idx = lane; idx < total; idx += 32) {
                int px, c8;
                split_idx(idx, p.cs8, sh_s, px, c8);
                if (px >= npx) continue;
                const float psi = s_psi[warp][px];
                *(reinterpret_cast<uint4*>(p.out + (opix0 + px) * p.out_ctot) + p.cx8 + c8) = scale_chunk<DT>(
                    __ldg(reinterpret_cast<const uint4*>(p.skip + (spix0 + px) * p.skip_ctot + p.skip_coff) + c8), psi);
            }
        }
        // ---- phase 4: slice 2 = spatially constant z_proj(z)
        {
            const int total = kCatPix * p.cz8;
            for (int idx = lane; idx < total; idx += 32) {
                int px, c8;
                split_idx(idx, p.cz8, sh_z, px, c8);
                if (px >= npx) continue;
                const float4* vp = reinterpret_cast<const float4*>(p.zvec + (size_t)d * p.cz8 * 8 + c8 * 8);
                const float4 a = __ldg(vp), b = __ldg(vp + 1);
                const float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
                *(reinterpret_cast<uint4*>(p.out + (opix0 + px) * p.out_ctot) + p.cx8 + p.cs8 + c8) = pack8(f, DT);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// fp16 fast path (the product dtype): same arithmetic per output element as decoder_concat_kernel<1>, reorganised so that
// the per-chunk instruction count drops ~5x (ncu r1e: the generic kernel is issue-bound at 2.8 TB/s, ~130 warp
// instructions per output pixel): a warp owns a run of 32 pixels of one output row and loops over the `rep` decodes of its
// patch (skip / x1 rows stay in L1 across the decodes); lanes keep a FIXED channel chunk and walk the pixels, so there is
// no index division, no per-chunk coordinate traffic through shared memory and all row base pointers are hoisted.
//   CX8 = chunks (8 channels) per pixel of the x slice and of the skip slice, F8 = chunks of the gate's F_int channels.
constexpr int kFastWarps = 4;
template <int CX8, int CS8, int F8> struct FastCat {
    static constexpr int RUN = 256 / CX8;          // pixels per warp run: 4 KB of the x slice per decode
    static constexpr int XP = 8;                   // x passes per run (RUN * CX8 / 32)
    static constexpr int SP = RUN * CS8 / 32;      // skip passes
    static constexpr int NG = 32 / F8;             // pixels per gate pass
    static constexpr int GP = RUN / NG;            // gate passes per run (<= F8)
    static constexpr int kWarpBytes = (RUN * CS8 + RUN * F8 + RUN) * 16 + RUN * 8 + RUN * 4;
    static constexpr int kSmemBytes = kFastWarps * kWarpBytes;
    static_assert(RUN >= 1 && RUN <= 32 && SP >= 1 && GP >= 1 && GP <= F8, "unsupported slice widths");
};

template <int CX8, int CS8, int F8>
__global__ void __launch_bounds__(kFastWarps * 32, 8)
decoder_concat_f16_kernel(const CatParams p, int runs_per_row, int num_runs) {
    using C = FastCat<CX8, CS8, F8>;
    constexpr int RUN = C::RUN, NG = C::NG, GP = C::GP;
    extern __shared__ uint4 smem_cat[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int run = blockIdx.x * kFastWarps + warp;
    if (run >= num_runs) return;       // warp-uniform; only __syncwarp below
    // per-warp staging: the run's skip and x1 pixels (shared by the `rep` decodes of the patch), bilinear weights / source
    // columns per pixel (same for all decodes), psi of the current decode
    uint4* s_skip = smem_cat + warp * (C::kWarpBytes / 16);
    uint4* s_x1 = s_skip + RUN * CS8;
    uint4* s_wt = s_x1 + RUN * F8;
    int2* s_off = reinterpret_cast<int2*>(s_wt + RUN);
    float* s_psi = reinterpret_cast<float*>(s_off + RUN);
    const int ho = run / runs_per_row;
    const int wo0 = (run - ho * runs_per_row) * RUN;
    const int npx = min(RUN, p.W - wo0);
    const uint32_t pb = blockIdx.y;
    // ATen upsample_bilinear2d, align_corners=True: src = dst * (in-1)/(out-1) in fp32
    const float h1r = p.sh * ho;
    const int h1 = (int)h1r;
    const int h1p = (h1 < p.hs - 1) ? 1 : 0;
    const float h1l = h1r - h1, h0l = 1.0f - h1l;
    const size_t spix0 = ((size_t)pb * p.H + ho) * p.W + wo0;            // skip / x1 pixel of px = 0
    const int gp = lane / F8, gf = lane % F8;
    {
        const uint16_t* skrow = p.skip + spix0 * p.skip_ctot + p.skip_coff;
#pragma unroll
        for (int pass = 0; pass < C::SP; ++pass) {
            const int idx = pass * 32 + lane, pxr = idx / CS8, c8 = idx % CS8;
            s_skip[idx] = __ldg(reinterpret_cast<const uint4*>(skrow + min(pxr, npx - 1) * p.skip_ctot) + c8);
        }
        const uint16_t* x1row = p.x1 + spix0 * p.x1_ctot + p.x1_coff + gf * 8;
#pragma unroll
        for (int pass = 0; pass < GP; ++pass) {
            const int pxr = pass * NG + gp;
            s_x1[pxr * F8 + gf] = __ldg(reinterpret_cast<const uint4*>(x1row + min(pxr, npx - 1) * p.x1_ctot));
        }
        if (lane < RUN) {
            const float w1r = p.sw * (wo0 + min(lane, npx - 1));
            const int w1 = min((int)w1r, p.ws - 1);
            const float w1l = w1r - (int)w1r, w0l = 1.0f - w1l;
            const __half2 w00 = __float2half2_rn(h0l * w0l), w01 = __float2half2_rn(h0l * w1l);
            const __half2 w10 = __float2half2_rn(h1l * w0l), w11 = __float2half2_rn(h1l * w1l);
            uint4 wt;
            wt.x = *reinterpret_cast<const uint32_t*>(&w00); wt.y = *reinterpret_cast<const uint32_t*>(&w01);
            wt.z = *reinterpret_cast<const uint32_t*>(&w10); wt.w = *reinterpret_cast<const uint32_t*>(&w11);
            s_wt[lane] = wt;
            s_off[lane] = make_int2(w1, w1 + ((w1 < p.ws - 1) ? 1 : 0));
        }
    }
    __syncwarp();
    auto lerp_tab = [&](const uint4& a, const uint4& b, const uint4& c, const uint4& d, const uint4& wt) {
        const __half2 w00 = *reinterpret_cast<const __half2*>(&wt.x), w01 = *reinterpret_cast<const __half2*>(&wt.y);
        const __half2 w10 = *reinterpret_cast<const __half2*>(&wt.z), w11 = *reinterpret_cast<const __half2*>(&wt.w);
        const __half2* pa = reinterpret_cast<const __half2*>(&a);
        const __half2* pb2 = reinterpret_cast<const __half2*>(&b);
        const __half2* pc = reinterpret_cast<const __half2*>(&c);
        const __half2* pd = reinterpret_cast<const __half2*>(&d);
        uint4 r;
        __half2* pr = reinterpret_cast<__half2*>(&r);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            pr[j] = __hfma2(w11, pd[j], __hfma2(w10, pc[j], __hfma2(w01, pb2[j], __hmul2(w00, pa[j]))));   // = lerp_chunk<1>
        return r;
    };
    float pw[8];
    {
        const float4 wa = __ldg(reinterpret_cast<const float4*>(p.psi_w) + gf * 2);
        const float4 wb = __ldg(reinterpret_cast<const float4*>(p.psi_w) + gf * 2 + 1);
        pw[0] = wa.x; pw[1] = wa.y; pw[2] = wa.z; pw[3] = wa.w; pw[4] = wb.x; pw[5] = wb.y; pw[6] = wb.z; pw[7] = wb.w;
    }
    const __half2 hzero = __float2half2_rn(0.0f);
    for (int r = 0; r < p.rep; ++r) {
        const size_t d = (size_t)pb * p.rep + r;
        const size_t srow0 = (d * p.hs + h1) * p.ws;                     // first source pixel of the upper source row
        const size_t srow1 = srow0 + (size_t)h1p * p.ws;
        const size_t opix0 = (d * p.H + ho) * p.W + wo0;
        // ---- phase 1: psi[px] = sigmoid(scale * <psi_w, relu(upsample(g1) + x1)> + bias). Pass `pass` covers pixels
        // pass*NG + gp; after the butterfly every lane of a pixel's group holds the dot product, and lane (gp, gf) keeps
        // the one of pass == gf: ONE sigmoid per lane (lanes with gf < GP) after the loop.
        {
            const uint16_t* g0 = p.g1 + p.g1_coff + srow0 * p.g1_ctot + gf * 8;
            const uint16_t* g1p = p.g1 + p.g1_coff + srow1 * p.g1_ctot + gf * 8;
            float mine = 0.0f;
#pragma unroll
            for (int pass = 0; pass < GP; ++pass) {
                const int px = pass * NG + gp;
                const int2 o = s_off[px];
                const uint4 gv = lerp_tab(__ldg(reinterpret_cast<const uint4*>(g0 + o.x * p.g1_ctot)),
                                          __ldg(reinterpret_cast<const uint4*>(g0 + o.y * p.g1_ctot)),
                                          __ldg(reinterpret_cast<const uint4*>(g1p + o.x * p.g1_ctot)),
                                          __ldg(reinterpret_cast<const uint4*>(g1p + o.y * p.g1_ctot)), s_wt[px]);
                const uint4 xv = s_x1[px * F8 + gf];
                const __half2* gh = reinterpret_cast<const __half2*>(&gv);
                const __half2* xh = reinterpret_cast<const __half2*>(&xv);
                float dsum = 0.0f;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 a = __half22float2(__hmax2(__hadd2(gh[j], xh[j]), hzero));
                    dsum = fmaf(pw[2 * j], a.x, dsum);
                    dsum = fmaf(pw[2 * j + 1], a.y, dsum);
                }
#pragma unroll
                for (int o2 = F8 >> 1; o2 > 0; o2 >>= 1) dsum += __shfl_xor_sync(0xffffffffu, dsum, o2);
                if (pass == gf) mine = dsum;
            }
            if (gf < GP) {
                const float psi = 1.0f / (1.0f + expf(-(p.psi_scale * mine + p.psi_bias)));
                const int pxm = gf * NG + gp;
                s_psi[pxm] = psi;
                if (p.psi_out != nullptr && pxm < npx) p.psi_out[opix0 + pxm] = psi;
            }
        }
        __syncwarp();
        // ---- phase 2: slice 0 = upsampled x
        {
            const uint16_t* x0 = p.x + p.x_coff + srow0 * p.x_ctot;
            const uint16_t* x1s = p.x + p.x_coff + srow1 * p.x_ctot;
            uint16_t* orow = p.out + opix0 * p.out_ctot;
#pragma unroll 4
            for (int pass = 0; pass < C::XP; ++pass) {
                const int idx = pass * 32 + lane, pxr = idx / CX8, c8 = idx % CX8;
                const int2 o = s_off[pxr];
                const uint4 v = lerp_tab(__ldg(reinterpret_cast<const uint4*>(x0 + o.x * p.x_ctot) + c8),
                                         __ldg(reinterpret_cast<const uint4*>(x0 + o.y * p.x_ctot) + c8),
                                         __ldg(reinterpret_cast<const uint4*>(x1s + o.x * p.x_ctot) + c8),
                                         __ldg(reinterpret_cast<const uint4*>(x1s + o.y * p.x_ctot) + c8), s_wt[pxr]);
                if (pxr < npx) *(reinterpret_cast<uint4*>(orow + pxr * p.out_ctot) + c8) = v;
            }
        }
        // ---- phase 3: slice 1 = gated skip
        {
            uint16_t* orow = p.out + opix0 * p.out_ctot + CX8 * 8;
#pragma unroll
            for (int pass = 0; pass < C::SP; ++pass) {
                const int idx = pass * 32 + lane, pxr = idx / CS8, c8 = idx % CS8;
                const uint4 v = scale_chunk<1>(s_skip[idx], s_psi[pxr]);
                if (pxr < npx) *(reinterpret_cast<uint4*>(orow + pxr * p.out_ctot) + c8) = v;
            }
        }
        __syncwarp();                  // s_psi is rewritten by the next decode
    }
}

template <int CX8, int CS8, int F8>
static void launch_concat_f16(const CatParams& p, cudaStream_t st) {
    using C = FastCat<CX8, CS8, F8>;
    static bool attr = false;
    if (!attr) {
        cudaFuncSetAttribute(decoder_concat_f16_kernel<CX8, CS8, F8>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes);
        attr = true;
    }
    const int runs_per_row = (p.W + C::RUN - 1) / C::RUN;
    const int num_runs = p.H * runs_per_row;
    const dim3 grid((num_runs + kFastWarps - 1) / kFastWarps, p.D / p.rep);
    decoder_concat_f16_kernel<CX8, CS8, F8><<<grid, kFastWarps * 32, C::kSmemBytes, st>>>(p, runs_per_row, num_runs);
}

// the slice widths of the reference's decoder blocks (in / skip / F_int = in/4: 512/256/128, 512/128/128, 256/64/64,
// 128/64/32, unet_resnet.py:228-231) + the square ones of the unit tests; anything else takes the generic kernel
static bool try_launch_concat_f16(const CatParams& p, cudaStream_t st) {
#define VU_CAT_CASE(CX, CS, F)                                              \
    if (p.cx8 == CX && p.cs8 == CS && p.f8 == F) {                          \
        launch_concat_f16<CX, CS, F>(p, st);                                \
        return true;                                                        \
    }
    VU_CAT_CASE(64, 32, 16)
    VU_CAT_CASE(64, 16, 16)
    VU_CAT_CASE(32, 8, 8)
    VU_CAT_CASE(16, 8, 4)
    VU_CAT_CASE(8, 8, 4)
    VU_CAT_CASE(16, 16, 8)
    VU_CAT_CASE(32, 32, 16)
#undef VU_CAT_CASE
    return false;
}

static int log2_exact(int v) {
    if (v <= 0 || (v & (v - 1))) return -1;
    int s = 0;
    while ((1 << s) < v) ++s;
    return s;
}

}  // namespace vu

using namespace vu;

extern "C" int vu_decoder_concat(const void* x_low, int hs, int ws, int Cx, int x_ctot, int x_coff, const void* g1_low,
                                 int g1_ctot, int g1_coff, const void* x1, int x1_ctot, int x1_coff, const void* skip,
                                 int skip_ctot, int skip_coff, int F, int Cs, const float* psi_w, float psi_scale,
                                 float psi_bias, const float* zvec, int Cz, void* out, int out_ctot, float* psi_out,
                                 int D, int rep, int H, int W, int dtype, vu_stream_t st) {
    if (!x_low || !g1_low || !x1 || !skip || !psi_w || !out) return set_error(VU_ERR_INVALID, "decoder_concat: null pointer");
    if (D < 1 || rep < 1 || D % rep || H < 1 || W < 1 || hs < 1 || ws < 1)
        return set_error(VU_ERR_INVALID, "decoder_concat: bad sizes (D=%d rep=%d H=%d W=%d hs=%d ws=%d)", D, rep, H, W, hs, ws);
    if ((Cx | Cs | F | Cz | x_ctot | x_coff | g1_ctot | g1_coff | x1_ctot | x1_coff | skip_ctot | skip_coff | out_ctot) % 8)
        return set_error(VU_ERR_ALIGN, "decoder_concat: channel counts/offsets must be multiples of 8");
    if (Cz > 0 && !zvec) return set_error(VU_ERR_INVALID, "decoder_concat: Cz > 0 without zvec");
    if (Cx + Cs + Cz > out_ctot) return set_error(VU_ERR_INVALID, "decoder_concat: slices exceed out_ctot");
    if ((size_t)D * H * W >= 0xffffffffull) return set_error(VU_ERR_UNSUPPORTED, "decoder_concat: more than 2^32 pixels");
    CatParams p{};
    p.x = (const uint16_t*)x_low; p.hs = hs; p.ws = ws; p.cx8 = Cx / 8; p.x_ctot = x_ctot; p.x_coff = x_coff;
    p.g1 = (const uint16_t*)g1_low; p.g1_ctot = g1_ctot; p.g1_coff = g1_coff;
    p.x1 = (const uint16_t*)x1; p.x1_ctot = x1_ctot; p.x1_coff = x1_coff;
    p.skip = (const uint16_t*)skip; p.skip_ctot = skip_ctot; p.skip_coff = skip_coff;
    p.f8 = F / 8; p.cs8 = Cs / 8;
    p.psi_w = psi_w; p.psi_scale = psi_scale; p.psi_bias = psi_bias;
    p.zvec = zvec; p.cz8 = Cz / 8;
    p.out = (uint16_t*)out; p.out_ctot = out_ctot; p.psi_out = psi_out;
    p.D = D; p.rep = rep; p.H = H; p.W = W;
    p.sh = H > 1 ? (float)(hs - 1) / (float)(H - 1) : 0.0f;
    p.sw = W > 1 ? (float)(ws - 1) / (float)(W - 1) : 0.0f;
    p.dtype = dtype;
    if (F > 128 || ((F / 8) & (F / 8 - 1)))
        return set_error(VU_ERR_UNSUPPORTED, "decoder_concat: F must be a power of two in [8, 128] (got %d)", F);
    const int segs_per_row = (W + kCatPix * kCatWarps - 1) / (kCatPix * kCatWarps);
    const size_t num_segs = (size_t)D * H * segs_per_row;
    size_t blocks = num_segs;
    const size_t cap = (size_t)num_sms() * 8;
    if (blocks > cap) blocks = cap;
    static const bool no_fast = getenv("VU_NO_FAST_CONCAT") != nullptr;
    const bool fast_ok = !no_fast && dtype == VU_F16 && Cz == 0 && D / rep <= 65535 && (size_t)ws * x_ctot < (1u << 30) &&
                         (size_t)ws * g1_ctot < (1u << 30) && (size_t)32 * out_ctot < (1u << 30);
    if (fast_ok && try_launch_concat_f16(p, (cudaStream_t)st)) {
    } else if (dtype == VU_F16)
        decoder_concat_kernel<1><<<(int)blocks, kCatWarps * 32, 0, (cudaStream_t)st>>>(
            p, segs_per_row, (uint32_t)num_segs, log2_exact(p.f8), log2_exact(p.cx8), log2_exact(p.cs8),
            p.cz8 ? log2_exact(p.cz8) : 0);
    else
        decoder_concat_kernel<0><<<(int)blocks, kCatWarps * 32, 0, (cudaStream_t)st>>>(
            p, segs_per_row, (uint32_t)num_segs, log2_exact(p.f8), log2_exact(p.cx8), log2_exact(p.cs8),
            p.cz8 ? log2_exact(p.cz8) : 0);
    VU_CHECK_LAUNCH("decoder_concat");
    return 0;
}
