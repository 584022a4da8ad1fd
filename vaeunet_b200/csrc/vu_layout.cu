// Memory-bound layout / pooling / resize kernels (NHWC 16-bit, 8 channels = 16 B per thread access).
#include "vu_internal.h"
#include <cstdlib>
#include "vu_half.cuh"

namespace vu {

// ---------------------------------------------------------------- weight packing
__global__ void pack_conv_weight_kernel(const float* __restrict__ w, uint16_t* __restrict__ out, int O, int I,
                                        int T, int transpose_io, int dtype) {
    const size_t total = (size_t)O * T * I;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int i = idx % I;
        const int t = (idx / I) % T;
        const int o = idx / ((size_t)I * T);
        const size_t src = transpose_io ? (((size_t)i * O + o) * T + t) : (((size_t)o * I + i) * T + t);
        out[idx] = f2h(w[src], dtype);
    }
}

// Filter banks of the DATA gradient of conv2d(x, w) (w OIHW, K x K, pad K/2), consumed by the forward conv kernels run on dY:
//   stride 1: one bank [I][K*K][O], taps rotated by 180 degrees;  stride 2, K = 1: one bank [I][1][O];
//   stride 2, K = 3: four 3x3 p1 banks [ph*2+pc][I][9][O], one per parity of the dX pixel (written by the pixel-shuffle
//   store): row parity 0 takes filter row r = 1 at bank row 1; parity 1 takes r = 2 at bank row 1 and r = 0 at bank row 2.
__device__ __forceinline__ int dgrad_s2_src(int parity, int bank_pos) {
    if (parity == 0) return bank_pos == 1 ? 1 : -1;
    return bank_pos == 1 ? 2 : (bank_pos == 2 ? 0 : -1);
}
__global__ void pack_dgrad_weight_kernel(const float* __restrict__ w, uint16_t* __restrict__ out, int O, int I, int K,
                                         int stride, int nbanks, int dtype) {
    const int T = K * K;
    const size_t per_bank = (size_t)I * T * O;
    const size_t total = per_bank * nbanks;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int o = idx % O;
        const int t = (idx / O) % T;
        const int i = (idx / ((size_t)O * T)) % I;
        const int bank = idx / per_bank;
        int r, s2;
        if (stride == 1) { r = K - 1 - t / K; s2 = K - 1 - t % K; }
        else if (K == 1) { r = 0; s2 = 0; }
        else { r = dgrad_s2_src(bank >> 1, t / 3); s2 = dgrad_s2_src(bank & 1, t % 3); }
        const float v = (r < 0 || s2 < 0) ? 0.f : w[(((size_t)o * I + i) * K + r) * K + s2];
        out[idx] = f2h(v, dtype);
    }
}

// rows of an fp32 matrix [O][I] -> 16-bit [O][I_pad], zero padded (the stem filter as a 1x1 bank over im2col patches)
__global__ void pack_rows_padded_kernel(const float* __restrict__ w, uint16_t* __restrict__ out, int O, int I, int I_pad, int dtype) {
    const size_t total = (size_t)O * I_pad;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int i = idx % I_pad, o = idx / I_pad;
        out[idx] = f2h(i < I ? w[(size_t)o * I + i] : 0.f, dtype);
    }
}

// im2col of a K x K stride-S pad-P convolution over an NCHW fp32 image with few channels (the ResNet stem, Cin = 3):
// out[n,ho,wo, ci*K*K + r*K + s] = x[n,ci,ho*S-P+r,wo*S-P+s] (0 outside the image and for the padding columns up to Kp).
// The patch matrix turns the stem conv and its weight gradient into 1x1 tensor-core GEMMs (training path).
// Block = 32 consecutive output pixels of one output row: the Cin x K x ((32-1)*S+K) input patch is staged in shared memory
// with coalesced loads, then every (pixel, 8-column chunk) is assembled from it and written as one 16-byte store (the
// 32 pixels are one contiguous run of the patch matrix).
constexpr int kI2cPix = 32;
__global__ void __launch_bounds__(256)
im2col_small_cin_kernel(const float* __restrict__ x, uint16_t* __restrict__ out, int N, int Cin, int H, int W, int K, int S,
                        int P, int Ho, int Wo, int Kp8, int dtype) {
    extern __shared__ float sm[];
    const int cols = (kI2cPix - 1) * S + K;
    float* tile = sm;                                        // [Cin][K][cols]
    int* tab = reinterpret_cast<int*>(sm + Cin * K * cols);  // [Kp8 * 8] smem offset of column kk (or -1)
    const int n = blockIdx.z, ho = blockIdx.y, wo0 = blockIdx.x * kI2cPix;
    const int KK = K * K, Ktot = Cin * KK;
    for (int i = threadIdx.x; i < Kp8 * 8; i += blockDim.x) {
        int off = -1;
        if (i < Ktot) { const int ci = i / KK, r = (i % KK) / K, q = i % K; off = (ci * K + r) * cols + q; }
        tab[i] = off;
    }
    for (int i = threadIdx.x; i < Cin * K * cols; i += blockDim.x) {
        const int c = i % cols, r = (i / cols) % K, ci = i / (cols * K);
        const int ih = ho * S - P + r, iw = wo0 * S - P + c;
        tile[i] = (ih >= 0 && ih < H && iw >= 0 && iw < W) ? __ldg(x + (((size_t)n * Cin + ci) * H + ih) * W + iw) : 0.f;
    }
    __syncthreads();
    const int npx = min(kI2cPix, Wo - wo0);
    uint4* dst = reinterpret_cast<uint4*>(out) + (((size_t)n * Ho + ho) * Wo + wo0) * Kp8;
    for (int i = threadIdx.x; i < npx * Kp8; i += blockDim.x) {
        const int px = i / Kp8, k8 = i % Kp8;
        float f[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int off = tab[k8 * 8 + j];
            f[j] = off >= 0 ? tile[off + px * S] : 0.f;
        }
        dst[i] = pack8(f, dtype);
    }
}

// Packed weight gradients [O][T][I] fp32 -> the parameters' own OIHW layout [O][I][T], ACCUMULATED, for a whole list of
// convolutions in one launch. table[e] = {src offset, dst offset (floats), O, T, I}; blockIdx.y = entry, blockIdx.x
// strides over the O rows; a row is a T x I matrix transposed through shared memory.
struct UnpackEntry { long long src, dst; int O, T, I, pad; };
__global__ void __launch_bounds__(256)
unpack_wgrad_batched_kernel(const float* __restrict__ src_base, float* __restrict__ dst_base,
                            const UnpackEntry* __restrict__ table) {
    extern __shared__ float row[];     // T rows of I + 4 floats: every I is a multiple of 32, so with a pitch of I the reads
                                       // row[t * I + ci] of 32 consecutive OIHW elements (9 taps x ~4 channels) fell on 4 banks;
                                       // with I + 4 they fall on 4 t + ci: all 32 banks
    const UnpackEntry e = table[blockIdx.y];
    const int n = e.T * e.I, pitch = e.I + 4;
    for (int o = blockIdx.x; o < e.O; o += gridDim.x) {
        const float* sp = src_base + e.src + (size_t)o * n;
        float* dp = dst_base + e.dst + (size_t)o * n;
        for (int t = 0; t < e.T; ++t)
            for (int ci = threadIdx.x; ci < e.I; ci += blockDim.x) row[t * pitch + ci] = sp[t * e.I + ci];
        __syncthreads();
        if (e.T == 9) {                                          // constant divisor: multiply-shift instead of a division
            for (int i = threadIdx.x; i < n; i += blockDim.x) {  // i = ci * T + t (OIHW order)
                const int ci = i / 9, t = i - ci * 9;
                dp[i] += row[t * pitch + ci];
            }
        } else if (e.T == 1) {
            for (int i = threadIdx.x; i < n; i += blockDim.x) dp[i] += row[i];
        } else {
            for (int i = threadIdx.x; i < n; i += blockDim.x) {
                const int ci = i / e.T, t = i % e.T;
                dp[i] += row[t * pitch + ci];
            }
        }
        __syncthreads();
    }
}

// Every tensor-core filter bank of a model in ONE launch (training repacks after each optimizer step): for table entry e
// the fp32 OIHW master at src -> the forward bank [O][K*K][I] at dst_fwd and the data-gradient bank(s) at dst_bank
// (same layouts as pack_conv_weight_kernel / pack_dgrad_weight_kernel). blockIdx.y = entry.
struct PackEntry { long long src, dst_fwd, dst_bank; int O, I, K, stride; };
// A block owns (entry, 32 output channels x 32 input channels): the 32 x (32*T) fp32 sub-matrix is read with coalesced
// rows into shared memory and written out twice - forward bank rows (32 consecutive i) and data-gradient bank rows (32
// consecutive o) - so both sides move 64..128-byte segments. blockIdx.y = entry, blockIdx.x strides over the tiles.
__global__ void __launch_bounds__(256)
pack_weights_batched_kernel(const float* __restrict__ src_base, uint16_t* __restrict__ dst_base,
                            const PackEntry* __restrict__ table, int dtype) {
    __shared__ float tile[32][32 * 9 + 1];
    const PackEntry e = table[blockIdx.y];
    const float* w = src_base + e.src;
    const int T = e.K * e.K, K = e.K;
    uint16_t* fwd = dst_base + e.dst_fwd;
    uint16_t* bank = dst_base + e.dst_bank;
    const int tiles_o = (e.O + 31) / 32, tiles_i = (e.I + 31) / 32;
    const int nbanks = (e.stride == 2 && K == 3) ? 4 : 1;
    const size_t per_bank = (size_t)e.I * T * e.O;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int tl = blockIdx.x; tl < tiles_o * tiles_i; tl += gridDim.x) {
        const int o0 = (tl / tiles_i) * 32, i0 = (tl % tiles_i) * 32;
        const int ni = min(32, e.I - i0), no = min(32, e.O - o0);
        const int rowlen = ni * T;
        __syncthreads();
        // no per-element div/mod anywhere: a warp takes whole rows, a lane a fixed (row, column) pair and walks the taps
        for (int oo = warp; oo < no; oo += 8) {
            const float* src = w + ((size_t)(o0 + oo) * e.I + i0) * T;      // ni*T contiguous floats
            for (int c = lane; c < rowlen; c += 32) tile[oo][c] = src[c];
        }
        __syncthreads();
        // forward bank [O][T][I]: lane = input channel (64-byte segments), warp strides over (output channel, tap)
        for (int q = warp; q < no * T; q += 8) {
            const int oo = q / T, t = q - oo * T;                           // once per 32 elements
            if (lane < ni) fwd[((size_t)(o0 + oo) * T + t) * e.I + i0 + lane] = f2h(tile[oo][lane * T + t], dtype);
        }
        // data-gradient bank(s) [b][I][T][O]: lane = output channel, warp strides over (bank, input channel, tap)
        for (int q = warp; q < nbanks * ni * T; q += 8) {
            const int bk = q / (ni * T), rem = q - bk * (ni * T);
            const int ii = rem / T, t = rem - ii * T;
            int r, s2;
            if (e.stride == 1) { r = K - 1 - t / K; s2 = K - 1 - t % K; }
            else if (K == 1) { r = 0; s2 = 0; }
            else { r = dgrad_s2_src(bk >> 1, t / 3); s2 = dgrad_s2_src(bk & 1, t % 3); }
            if (lane < no) {
                const float v = (r < 0 || s2 < 0) ? 0.f : tile[lane][ii * T + r * K + s2];
                bank[bk * per_bank + ((size_t)(i0 + ii) * T + t) * e.O + o0 + lane] = f2h(v, dtype);
            }
        }
    }
}

// ConvTranspose2d k2 s2: w [I][O][2][2] -> [4 (a*2+b)][O][I]
__global__ void pack_convT_kernel(const float* __restrict__ w, uint16_t* __restrict__ out, int I, int O, int dtype) {
    const size_t total = (size_t)4 * O * I;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int i = idx % I;
        const int o = (idx / I) % O;
        const int ab = idx / ((size_t)I * O);
        out[idx] = f2h(w[((size_t)i * O + o) * 4 + ab], dtype);
    }
}

// ---------------------------------------------------------------- NCHW fp32 <-> NHWC 16-bit
// 32 pixels x 32 channels tile through smem: coalesced on both sides.
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ x, uint16_t* __restrict__ y, int C, int HW,
                                    int y_ctot, int y_coff, int dtype) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z;
    const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;  // 32 x 8
    for (int j = ty; j < 32; j += 8) {
        const int c = c0 + j, p = p0 + tx;
        tile[j][tx] = (c < C && p < HW) ? x[((size_t)n * C + c) * HW + p] : 0.0f;
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        const int p = p0 + j, c = c0 + tx;
        if (c < C && p < HW) y[((size_t)n * HW + p) * y_ctot + y_coff + c] = f2h(tile[tx][j], dtype);
    }
}

__global__ void nhwc_to_nchw_kernel(const uint16_t* __restrict__ x, float* __restrict__ y, int C, int HW,
                                    int x_ctot, int x_coff, int dtype) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z;
    const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
    for (int j = ty; j < 32; j += 8) {
        const int p = p0 + j, c = c0 + tx;
        tile[j][tx] = (c < C && p < HW) ? h2f(x[((size_t)n * HW + p) * x_ctot + x_coff + c], dtype) : 0.0f;
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        const int c = c0 + j, p = p0 + tx;
        if (c < C && p < HW) y[((size_t)n * C + c) * HW + p] = tile[tx][j];
    }
}

// ---------------------------------------------------------------- direct small-Cin convs
// Generic direct conv for Cin <= 4 (stem 7x7 s2 p3; basic-UNet inc 3x3 s1 p1), fp32 weights and pixels. Persistent CTAs:
// the whole filter bank is staged in shared memory ONCE per CTA (coalesced read, tap-major so a warp reads one row as a
// broadcast), then the CTA walks over 16 x 32 output tiles; a thread owns two pixels of a row (16 columns apart) x 32 output
// channels per pass, so every 16-byte weight read feeds 8 FMAs.
template <int K, int S>
__global__ void __launch_bounds__(256)
small_cin_conv_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ scale,
                      const float* __restrict__ bias, uint16_t* __restrict__ y, int N, int Cin, int H, int W, int Cout,
                      int Ho, int Wo, int pad, int y_ctot, int y_coff, int relu, int dtype) {
    constexpr int TOH = 16, TOW = 32;
    constexpr int TIH = (TOH - 1) * S + K, TIW = (TOW - 1) * S + K;
    extern __shared__ float sm[];
    float* s_in = sm;                              // [Cin][TIH][TIW+1]
    float* s_w = sm + 4 * TIH * (TIW + 1);         // [Cin*K*K][Cout]
    const int tid = threadIdx.x;
    const int KK = Cin * K * K;
    for (int i = tid; i < KK * Cout; i += 256) {   // i walks the OIHW tensor linearly (coalesced)
        const int co = i / KK, t = i % KK;
        s_w[t * Cout + co] = w[i];
    }
    const int tiles_w = (Wo + TOW - 1) / TOW, tiles_h = (Ho + TOH - 1) / TOH;
    const int total = tiles_w * tiles_h * N;
    const int ty = tid / 16, tx = tid % 16;
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int n = tile / (tiles_w * tiles_h);
        const int oh0 = ((tile / tiles_w) % tiles_h) * TOH, ow0 = (tile % tiles_w) * TOW;
        const int ih0 = oh0 * S - pad, iw0 = ow0 * S - pad;
        __syncthreads();
        for (int i = tid; i < Cin * TIH * TIW; i += 256) {
            const int c = i / (TIH * TIW), r = (i / TIW) % TIH, q = i % TIW;
            const int ih = ih0 + r, iw = iw0 + q;
            float v = 0.0f;
            if (ih >= 0 && ih < H && iw >= 0 && iw < W) v = __ldg(x + (((size_t)n * Cin + c) * H + ih) * W + iw);
            s_in[(c * TIH + r) * (TIW + 1) + q] = v;
        }
        __syncthreads();
        const int oh = oh0 + ty;
        for (int cb = 0; cb < Cout; cb += 32) {
            float acc[2][32];
#pragma unroll
            for (int j = 0; j < 32; ++j) { acc[0][j] = 0.0f; acc[1][j] = 0.0f; }
            for (int c = 0; c < Cin; ++c)
                for (int r = 0; r < K; ++r) {
                    const float* row = s_in + (c * TIH + ty * S + r) * (TIW + 1) + tx * S;
#pragma unroll
                    for (int q = 0; q < K; ++q) {
                        const float v0 = row[q], v1 = row[q + 16 * S];
                        const float4* wp = reinterpret_cast<const float4*>(s_w + ((c * K + r) * K + q) * Cout + cb);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 ww = wp[j];
                            acc[0][4 * j] = fmaf(v0, ww.x, acc[0][4 * j]);         acc[1][4 * j] = fmaf(v1, ww.x, acc[1][4 * j]);
                            acc[0][4 * j + 1] = fmaf(v0, ww.y, acc[0][4 * j + 1]); acc[1][4 * j + 1] = fmaf(v1, ww.y, acc[1][4 * j + 1]);
                            acc[0][4 * j + 2] = fmaf(v0, ww.z, acc[0][4 * j + 2]); acc[1][4 * j + 2] = fmaf(v1, ww.z, acc[1][4 * j + 2]);
                            acc[0][4 * j + 3] = fmaf(v0, ww.w, acc[0][4 * j + 3]); acc[1][4 * j + 3] = fmaf(v1, ww.w, acc[1][4 * j + 3]);
                        }
                    }
                }
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int ow = ow0 + tx + 16 * h;
                if (oh < Ho && ow < Wo) {
                    uint32_t o[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        float a = acc[h][2 * j] * scale[cb + 2 * j] + bias[cb + 2 * j];
                        float b = acc[h][2 * j + 1] * scale[cb + 2 * j + 1] + bias[cb + 2 * j + 1];
                        if (relu) { a = fmaxf(a, 0.0f); b = fmaxf(b, 0.0f); }
                        o[j] = pack2h(a, b, dtype);
                    }
                    uint4* op = reinterpret_cast<uint4*>(y + (((size_t)n * Ho + oh) * Wo + ow) * y_ctot + y_coff + cb);
#pragma unroll
                    for (int j = 0; j < 4; ++j) op[j] = make_uint4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
                }
            }
        }
    }
}

template <int K, int S>
static int launch_small_cin(const float* x, const float* w, const float* scale, const float* bias, void* y, int N,
                            int Cin, int H, int W, int Cout, int pad, int y_ctot, int y_coff, int relu, int dtype,
                            cudaStream_t stream) {
    constexpr int TIH = 15 * S + K, TIW = 31 * S + K;
    if (Cin < 1 || Cin > 4) return set_error(VU_ERR_UNSUPPORTED, "small-cin conv: Cin must be 1..4 (got %d)", Cin);
    if (Cout % 32 || y_ctot % 8 || y_coff % 8) return set_error(VU_ERR_ALIGN, "small-cin conv: Cout %% 32, slices %% 8");
    if (!scale || !bias) return set_error(VU_ERR_INVALID, "small-cin conv: scale/bias required");
    static const bool fma_only = getenv("VU_STEM_FMA") != nullptr;      // A/B switch: the FP32-FMA direct convolution
    if (!fma_only && launch_small_cin_mma(K, S, x, w, scale, bias, y, N, Cin, H, W, Cout, pad, y_ctot, y_coff, relu, dtype,
                                          stream) == 0) {
        VU_CHECK_LAUNCH("small_cin_mma");
        return 0;
    }
    const int Ho = (H + 2 * pad - K) / S + 1, Wo = (W + 2 * pad - K) / S + 1;
    const size_t smem = (size_t)(4 * TIH * (TIW + 1) + Cin * K * K * Cout) * sizeof(float);
    static bool attr = false;
    if (!attr) {
        cudaFuncSetAttribute(small_cin_conv_kernel<K, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        attr = true;
    }
    if (smem > 200 * 1024) return set_error(VU_ERR_UNSUPPORTED, "small-cin conv: filter bank too large for smem");
    const int total = ((Wo + 31) / 32) * ((Ho + 15) / 16) * N;
    const int per_sm = (int)((220 * 1024) / (smem + 1024));
    int grid = num_sms() * (per_sm < 1 ? 1 : (per_sm > 3 ? 3 : per_sm));
    if (grid > total) grid = total;
    small_cin_conv_kernel<K, S><<<grid, 256, smem, stream>>>(x, w, scale, bias, static_cast<uint16_t*>(y), N, Cin, H, W,
                                                           Cout, Ho, Wo, pad, y_ctot, y_coff, relu, dtype);
    VU_CHECK_LAUNCH("small_cin_conv");
    return 0;
}

// ---------------------------------------------------------------- max-pool
__global__ void maxpool_nhwc_kernel(const uint16_t* __restrict__ x, uint16_t* __restrict__ y, int N, int H, int W,
                                    int C8, int x_ctot, int x_coff, int k, int s, int p, int Ho, int Wo, int y_ctot,
                                    int y_coff, int dtype) {
    const size_t total = (size_t)N * Ho * Wo * C8;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int c8 = idx % C8;
        const int wo = (idx / C8) % Wo;
        const int ho = (idx / ((size_t)C8 * Wo)) % Ho;
        const int n = idx / ((size_t)C8 * Wo * Ho);
        float m[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) m[j] = -INFINITY;
        for (int r = 0; r < k; ++r) {
            const int ih = ho * s - p + r;
            if (ih < 0 || ih >= H) continue;
            for (int q = 0; q < k; ++q) {
                const int iw = wo * s - p + q;
                if (iw < 0 || iw >= W) continue;
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(x + (((size_t)n * H + ih) * W + iw) * x_ctot + x_coff) + c8);
                float f[8];
                unpack8(v, f, dtype);
#pragma unroll
                for (int j = 0; j < 8; ++j) m[j] = fmaxf(m[j], f[j]);
            }
        }
        *(reinterpret_cast<uint4*>(y + (((size_t)n * Ho + ho) * Wo + wo) * y_ctot + y_coff) + c8) = pack8(m, dtype);
    }
}

// ---------------------------------------------------------------- bilinear, align_corners=True
// Coordinate arithmetic follows ATen's upsample_bilinear2d (area_pixel_compute_scale /
// _source_index with align_corners): scale = (in-1)/(out-1) in fp32, src = scale*dst.
// I = index type: 32-bit arithmetic whenever the element count allows (64-bit div/mod costs more than the memory traffic)
template <typename I>
__global__ void __launch_bounds__(256) bilinear_ac_nhwc_kernel(const uint16_t* __restrict__ x, uint16_t* __restrict__ y, int N, int Hs, int Ws,
                                        int C8, int x_ctot, int x_coff, int Ho, int Wo, int Hd, int Wd, int pad_t,
                                        int pad_l, int y_ctot, int y_coff, float sh, float sw, int dtype) {
    const I total = (I)N * Ho * Wo * C8;
    for (I idx = blockIdx.x * (I)blockDim.x + threadIdx.x; idx < total; idx += (I)gridDim.x * blockDim.x) {
        const int c8 = idx % C8;
        const int wo = (idx / C8) % Wo;
        const int ho = (idx / ((I)C8 * Wo)) % Ho;
        const int n = idx / ((I)C8 * Wo * Ho);
        const float h1r = sh * ho;
        const int h1 = (int)h1r;
        const int h1p = (h1 < Hs - 1) ? 1 : 0;
        const float h1l = h1r - h1, h0l = 1.0f - h1l;
        const float w1r = sw * wo;
        const int w1 = (int)w1r;
        const int w1p = (w1 < Ws - 1) ? 1 : 0;
        const float w1l = w1r - w1, w0l = 1.0f - w1l;
        const uint16_t* base = x + ((size_t)n * Hs * Ws) * x_ctot + x_coff;
        const uint4 v00 = __ldg(reinterpret_cast<const uint4*>(base + ((size_t)h1 * Ws + w1) * x_ctot) + c8);
        const uint4 v01 = __ldg(reinterpret_cast<const uint4*>(base + ((size_t)h1 * Ws + w1 + w1p) * x_ctot) + c8);
        const uint4 v10 = __ldg(reinterpret_cast<const uint4*>(base + ((size_t)(h1 + h1p) * Ws + w1) * x_ctot) + c8);
        const uint4 v11 = __ldg(reinterpret_cast<const uint4*>(base + ((size_t)(h1 + h1p) * Ws + w1 + w1p) * x_ctot) + c8);
        float a[8], b[8], c[8], d[8], o[8];
        unpack8(v00, a, dtype); unpack8(v01, b, dtype); unpack8(v10, c, dtype); unpack8(v11, d, dtype);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = h0l * (w0l * a[j] + w1l * b[j]) + h1l * (w0l * c[j] + w1l * d[j]);
        *(reinterpret_cast<uint4*>(y + (((size_t)n * Hd + ho + pad_t) * Wd + wo + pad_l) * y_ctot + y_coff) + c8) =
            pack8(o, dtype);
    }
}

// Fast path of the kernel above for power-of-two channel counts (C8 = 4..64 chunks per pixel): a warp owns a run of
// RUN = min(32, 256 / C8) pixels of one output row, the per-pixel source columns and column weights go through a per-warp
// shared-memory table (computed once per run instead of once per 16-byte chunk), lanes walk the flattened (pixel, chunk)
// list of the run with compile-time shifts - no integer division, one row base pointer per run. Arithmetic (fp32, ATen's
// order) is the generic kernel's, results are bit-identical. ncu of the generic kernel: ~250 instructions per chunk (four
// div/mod pairs, 64-bit addressing), issue-bound at ~40 % of the write bandwidth.
template <int C8>
__global__ void __launch_bounds__(128)
bilinear_ac_run_kernel(const uint16_t* __restrict__ x, uint16_t* __restrict__ y, int Hs, int Ws, int x_ctot, int x_coff, int Ho,
                       int Wo, int Hd, int Wd, int pad_t, int pad_l, int y_ctot, int y_coff, float sh, float sw, int dtype,
                       int runs_per_row, int num_runs) {
    constexpr int RUN = (256 / C8 < 32) ? 256 / C8 : 32;
    constexpr int PASSES = RUN * C8 / 32;
    __shared__ int2 s_off[4][32];
    __shared__ float s_w1l[4][32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int run = blockIdx.x * 4 + warp;
    if (run >= num_runs) return;                 // warp-uniform; only __syncwarp below
    const int n = blockIdx.y;
    const int ho = run / runs_per_row;
    const int wo0 = (run - ho * runs_per_row) * RUN;
    const int npx = min(RUN, Wo - wo0);
    const float h1r = sh * ho;
    const int h1 = (int)h1r;
    const int h1p = (h1 < Hs - 1) ? 1 : 0;
    const float h1l = h1r - h1, h0l = 1.0f - h1l;
    if (lane < RUN) {
        const float w1r = sw * (wo0 + min(lane, npx - 1));
        const int w1 = (int)w1r;
        s_off[warp][lane] = make_int2(w1, w1 + ((w1 < Ws - 1) ? 1 : 0));
        s_w1l[warp][lane] = w1r - w1;
    }
    __syncwarp();
    const uint16_t* r0 = x + ((size_t)n * Hs + h1) * Ws * x_ctot + x_coff;
    const uint16_t* r1 = r0 + (size_t)h1p * Ws * x_ctot;
    uint16_t* orow = y + (((size_t)n * Hd + ho + pad_t) * Wd + wo0 + pad_l) * y_ctot + y_coff;
#pragma unroll 4
    for (int pass = 0; pass < PASSES; ++pass) {
        const int idx = pass * 32 + lane, px = idx / C8, c8 = idx % C8;
        const int2 o = s_off[warp][px];
        const float w1l = s_w1l[warp][px], w0l = 1.0f - w1l;
        const uint4 v00 = __ldg(reinterpret_cast<const uint4*>(r0 + o.x * x_ctot) + c8);
        const uint4 v01 = __ldg(reinterpret_cast<const uint4*>(r0 + o.y * x_ctot) + c8);
        const uint4 v10 = __ldg(reinterpret_cast<const uint4*>(r1 + o.x * x_ctot) + c8);
        const uint4 v11 = __ldg(reinterpret_cast<const uint4*>(r1 + o.y * x_ctot) + c8);
        float a[8], b[8], c[8], d[8], q[8];
        unpack8(v00, a, dtype); unpack8(v01, b, dtype); unpack8(v10, c, dtype); unpack8(v11, d, dtype);
#pragma unroll
        for (int j = 0; j < 8; ++j) q[j] = h0l * (w0l * a[j] + w1l * b[j]) + h1l * (w0l * c[j] + w1l * d[j]);
        if (px < npx) *(reinterpret_cast<uint4*>(orow + px * y_ctot) + c8) = pack8(q, dtype);
    }
}

template <int C8>
static void launch_bilinear_run(const uint16_t* x, uint16_t* y, int N, int Hs, int Ws, int x_ctot, int x_coff, int Ho, int Wo,
                                int Hd, int Wd, int pad_t, int pad_l, int y_ctot, int y_coff, float sh, float sw, int dtype,
                                cudaStream_t st) {
    constexpr int RUN = (256 / C8 < 32) ? 256 / C8 : 32;
    const int runs_per_row = (Wo + RUN - 1) / RUN, num_runs = Ho * runs_per_row;
    bilinear_ac_run_kernel<C8><<<dim3((num_runs + 3) / 4, N), 128, 0, st>>>(x, y, Hs, Ws, x_ctot, x_coff, Ho, Wo, Hd, Wd, pad_t,
                                                                          pad_l, y_ctot, y_coff, sh, sw, dtype, runs_per_row,
                                                                          num_runs);
}

__global__ void broadcast_vec_nhwc_kernel(const float* __restrict__ v, uint16_t* __restrict__ y, int N, int HW, int C8,
                                          int C, int y_ctot, int y_coff, int dtype) {
    const size_t total = (size_t)N * HW * C8;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int c8 = idx % C8;
        const size_t pix = idx / C8;
        const int n = pix / HW;
        const float4* vp = reinterpret_cast<const float4*>(v + (size_t)n * C + c8 * 8);
        const float4 a = __ldg(vp), b = __ldg(vp + 1);
        const float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        *(reinterpret_cast<uint4*>(y + pix * y_ctot + y_coff) + c8) = pack8(f, dtype);
    }
}

static int grid_for(size_t total, int block) {
    size_t g = (total + block - 1) / block;
    const size_t cap = (size_t)num_sms() * 16;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

}  // namespace vu

using namespace vu;

extern "C" int vu_pack_conv_weight(const float* w, void* out, int O, int I, int KH, int KW, int transpose_io, int dtype,
                                   vu_stream_t st) {
    if (!w || !out || O < 1 || I < 1) return set_error(VU_ERR_INVALID, "pack_conv_weight: bad args");
    const size_t total = (size_t)O * I * KH * KW;
    pack_conv_weight_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)st>>>(w, (uint16_t*)out, O, I, KH * KW,
                                                                               transpose_io, dtype);
    VU_CHECK_LAUNCH("pack_conv_weight");
    return 0;
}

extern "C" int vu_unpack_wgrad_batched(const float* packed_base, float* oihw_base, const void* table, int n_entries,
                                       int max_row_floats, vu_stream_t st) {
    if (!packed_base || !oihw_base || !table || n_entries < 1 || max_row_floats < 1)
        return set_error(VU_ERR_INVALID, "unpack_wgrad_batched: bad args");
    const size_t smem = ((size_t)max_row_floats + 256) * sizeof(float);       // + four pad floats per tap row (T <= 64)
    if (smem > 160 * 1024) return set_error(VU_ERR_UNSUPPORTED, "unpack_wgrad_batched: row of %d floats exceeds shared memory", max_row_floats);
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(unpack_wgrad_batched_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
        attr_set = true;
    }
    dim3 grid(128, n_entries);
    unpack_wgrad_batched_kernel<<<grid, 256, smem, (cudaStream_t)st>>>(packed_base, oihw_base, (const UnpackEntry*)table);
    VU_CHECK_LAUNCH("unpack_wgrad_batched");
    return 0;
}

extern "C" int vu_pack_weights_batched(const float* src_base, void* dst_base, const void* table, int n_entries, int dtype,
                                       vu_stream_t st) {
    if (!src_base || !dst_base || !table || n_entries < 1) return set_error(VU_ERR_INVALID, "pack_weights_batched: bad args");
    dim3 grid(416, n_entries);      // >= the tile count of the largest bank (800 x 512): one tile per block, empty blocks exit
                                    // (104 blocks striding four rounds measured slower: 0.28 vs 0.24 ms)
    pack_weights_batched_kernel<<<grid, 256, 0, (cudaStream_t)st>>>(src_base, (uint16_t*)dst_base, (const PackEntry*)table, dtype);
    VU_CHECK_LAUNCH("pack_weights_batched");
    return 0;
}

extern "C" int vu_pack_rows_padded(const float* w, void* out, int O, int I, int I_pad, int dtype, vu_stream_t st) {
    if (!w || !out || O < 1 || I < 1 || I_pad < I) return set_error(VU_ERR_INVALID, "pack_rows_padded: bad args");
    pack_rows_padded_kernel<<<grid_for((size_t)O * I_pad, 256), 256, 0, (cudaStream_t)st>>>(w, (uint16_t*)out, O, I, I_pad, dtype);
    VU_CHECK_LAUNCH("pack_rows_padded");
    return 0;
}

extern "C" int vu_im2col_small_cin(const float* x_nchw, void* out, int N, int Cin, int H, int W, int K, int S, int P, int Kp,
                                   int dtype, vu_stream_t st) {
    if (!x_nchw || !out || Cin < 1 || K < 1 || S < 1 || Kp % 8 || Kp < Cin * K * K)
        return set_error(VU_ERR_INVALID, "im2col_small_cin: bad args (Kp multiple of 8, >= Cin*K*K)");
    const int Ho = (H + 2 * P - K) / S + 1, Wo = (W + 2 * P - K) / S + 1;
    if (Cin > 4 || K > 7 || S > 2) return set_error(VU_ERR_UNSUPPORTED, "im2col_small_cin: Cin <= 4, K <= 7, S <= 2");
    const int cols = (kI2cPix - 1) * S + K;
    const size_t smem = (size_t)Cin * K * cols * sizeof(float) + (size_t)Kp * sizeof(int);
    dim3 grid((Wo + kI2cPix - 1) / kI2cPix, Ho, N);
    im2col_small_cin_kernel<<<grid, 256, smem, (cudaStream_t)st>>>(x_nchw, (uint16_t*)out, N, Cin, H, W, K, S, P, Ho, Wo,
                                                                  Kp / 8, dtype);
    VU_CHECK_LAUNCH("im2col_small_cin");
    return 0;
}

extern "C" int vu_pack_dgrad_weight(const float* w, void* banks, int O, int I, int K, int stride, int dtype, vu_stream_t st) {
    if (!w || !banks || O < 1 || I < 1) return set_error(VU_ERR_INVALID, "pack_dgrad_weight: bad args");
    if (!((K == 3 || K == 1) && (stride == 1 || stride == 2)))
        return set_error(VU_ERR_UNSUPPORTED, "pack_dgrad_weight: only 3x3 / 1x1, stride 1|2 (got k%d s%d)", K, stride);
    const int nbanks = (stride == 2 && K == 3) ? 4 : 1;
    const size_t total = (size_t)nbanks * O * I * K * K;
    pack_dgrad_weight_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)st>>>(w, (uint16_t*)banks, O, I, K, stride, nbanks, dtype);
    VU_CHECK_LAUNCH("pack_dgrad_weight");
    return 0;
}

extern "C" int vu_pack_convT2x2_weight(const float* w, void* out, int I, int O, int dtype, vu_stream_t st) {
    if (!w || !out) return set_error(VU_ERR_INVALID, "pack_convT: null");
    pack_convT_kernel<<<grid_for((size_t)4 * I * O, 256), 256, 0, (cudaStream_t)st>>>(w, (uint16_t*)out, I, O, dtype);
    VU_CHECK_LAUNCH("pack_convT");
    return 0;
}

extern "C" int vu_nchw_to_nhwc(const float* x, void* y, int N, int C, int H, int W, int y_ctot, int y_coff, int dtype,
                               vu_stream_t st) {
    if (!x || !y || N < 1 || C < 1) return set_error(VU_ERR_INVALID, "nchw_to_nhwc: bad args");
    dim3 grid((H * W + 31) / 32, (C + 31) / 32, N), block(32, 8);
    nchw_to_nhwc_kernel<<<grid, block, 0, (cudaStream_t)st>>>(x, (uint16_t*)y, C, H * W, y_ctot, y_coff, dtype);
    VU_CHECK_LAUNCH("nchw_to_nhwc");
    return 0;
}

extern "C" int vu_nhwc_to_nchw(const void* x, float* y, int N, int C, int H, int W, int x_ctot, int x_coff, int dtype,
                               vu_stream_t st) {
    if (!x || !y || N < 1 || C < 1) return set_error(VU_ERR_INVALID, "nhwc_to_nchw: bad args");
    dim3 grid((H * W + 31) / 32, (C + 31) / 32, N), block(32, 8);
    nhwc_to_nchw_kernel<<<grid, block, 0, (cudaStream_t)st>>>((const uint16_t*)x, y, C, H * W, x_ctot, x_coff, dtype);
    VU_CHECK_LAUNCH("nhwc_to_nchw");
    return 0;
}

extern "C" int vu_stem_conv7x7s2(const float* x, const float* w, const float* scale, const float* bias, void* y, int N,
                                 int Cin, int H, int W, int Cout, int y_ctot, int y_coff, int dtype, vu_stream_t st) {
    if (!x || !w || !y) return set_error(VU_ERR_INVALID, "stem: null");
    return launch_small_cin<7, 2>(x, w, scale, bias, y, N, Cin, H, W, Cout, 3, y_ctot, y_coff, 1, dtype, (cudaStream_t)st);
}

extern "C" int vu_stem_conv7x7s2_act(const float* x, const float* w, const float* scale, const float* bias, void* y, int N,
                                     int Cin, int H, int W, int Cout, int y_ctot, int y_coff, int relu, int dtype,
                                     vu_stream_t st) {
    if (!x || !w || !y) return set_error(VU_ERR_INVALID, "stem: null");
    return launch_small_cin<7, 2>(x, w, scale, bias, y, N, Cin, H, W, Cout, 3, y_ctot, y_coff, relu, dtype, (cudaStream_t)st);
}

extern "C" int vu_conv3x3_small_cin(const float* x, const float* w, const float* scale, const float* bias, void* y,
                                    int N, int Cin, int H, int W, int Cout, int y_ctot, int y_coff, int relu, int dtype,
                                    vu_stream_t st) {
    if (!x || !w || !y) return set_error(VU_ERR_INVALID, "conv3x3_small_cin: null");
    return launch_small_cin<3, 1>(x, w, scale, bias, y, N, Cin, H, W, Cout, 1, y_ctot, y_coff, relu, dtype, (cudaStream_t)st);
}

extern "C" int vu_maxpool_nhwc(const void* x, void* y, int N, int H, int W, int C, int x_ctot, int x_coff, int k, int s,
                               int p, int y_ctot, int y_coff, int dtype, vu_stream_t st) {
    if (!x || !y) return set_error(VU_ERR_INVALID, "maxpool: null");
    if (C % 8 || x_ctot % 8 || x_coff % 8 || y_ctot % 8 || y_coff % 8) return set_error(VU_ERR_ALIGN, "maxpool: channels %% 8");
    const int Ho = (H + 2 * p - k) / s + 1, Wo = (W + 2 * p - k) / s + 1;
    const size_t total = (size_t)N * Ho * Wo * (C / 8);
    maxpool_nhwc_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)st>>>((const uint16_t*)x, (uint16_t*)y, N, H, W, C / 8,
                                                                           x_ctot, x_coff, k, s, p, Ho, Wo, y_ctot, y_coff, dtype);
    VU_CHECK_LAUNCH("maxpool");
    return 0;
}

extern "C" int vu_bilinear_ac_nhwc(const void* x, void* y, int N, int Hs, int Ws, int C, int x_ctot, int x_coff, int Ho,
                                   int Wo, int Hd, int Wd, int pad_t, int pad_l, int y_ctot, int y_coff, int dtype,
                                   vu_stream_t st) {
    if (!x || !y) return set_error(VU_ERR_INVALID, "bilinear: null");
    if (C % 8 || x_ctot % 8 || x_coff % 8 || y_ctot % 8 || y_coff % 8) return set_error(VU_ERR_ALIGN, "bilinear: channels %% 8");
    if (pad_t < 0 || pad_l < 0 || pad_t + Ho > Hd || pad_l + Wo > Wd) return set_error(VU_ERR_INVALID, "bilinear: bad padding");
    const float sh = Ho > 1 ? (float)(Hs - 1) / (float)(Ho - 1) : 0.0f;
    const float sw = Wo > 1 ? (float)(Ws - 1) / (float)(Wo - 1) : 0.0f;
    const size_t total = (size_t)N * Ho * Wo * (C / 8);
    static const bool generic_only = getenv("VU_BILINEAR_GENERIC") != nullptr;
    const bool run_ok = !generic_only && N <= 65535 && (size_t)Ws * x_ctot < (1u << 30) && (size_t)32 * y_ctot < (1u << 30);
#define VU_BIL_CASE(c8)                                                                                                         \
    if (run_ok && C / 8 == c8) {                                                                                                \
        launch_bilinear_run<c8>((const uint16_t*)x, (uint16_t*)y, N, Hs, Ws, x_ctot, x_coff, Ho, Wo, Hd, Wd, pad_t, pad_l, y_ctot, \
                                y_coff, sh, sw, dtype, (cudaStream_t)st);                                                       \
        VU_CHECK_LAUNCH("bilinear_run");                                                                                        \
        return 0;                                                                                                               \
    }
    VU_BIL_CASE(4) VU_BIL_CASE(8) VU_BIL_CASE(16) VU_BIL_CASE(32) VU_BIL_CASE(64)
#undef VU_BIL_CASE
    if (total < (1ull << 31))
        bilinear_ac_nhwc_kernel<uint32_t><<<grid_for(total, 256), 256, 0, (cudaStream_t)st>>>(
            (const uint16_t*)x, (uint16_t*)y, N, Hs, Ws, C / 8, x_ctot, x_coff, Ho, Wo, Hd, Wd, pad_t, pad_l, y_ctot, y_coff, sh, sw, dtype);
    else
        bilinear_ac_nhwc_kernel<size_t><<<grid_for(total, 256), 256, 0, (cudaStream_t)st>>>(
            (const uint16_t*)x, (uint16_t*)y, N, Hs, Ws, C / 8, x_ctot, x_coff, Ho, Wo, Hd, Wd, pad_t, pad_l, y_ctot, y_coff, sh, sw, dtype);
    VU_CHECK_LAUNCH("bilinear");
    return 0;
}

extern "C" int vu_broadcast_vec_nhwc(const float* v, void* y, int N, int H, int W, int C, int y_ctot, int y_coff, int dtype,
                                     vu_stream_t st) {
    if (!v || !y) return set_error(VU_ERR_INVALID, "broadcast: null");
    if (C % 8 || y_ctot % 8 || y_coff % 8) return set_error(VU_ERR_ALIGN, "broadcast: channels %% 8");
    const size_t total = (size_t)N * H * W * (C / 8);
    broadcast_vec_nhwc_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)st>>>(v, (uint16_t*)y, N, H * W, C / 8, C, y_ctot,
                                                                                 y_coff, dtype);
    VU_CHECK_LAUNCH("broadcast");
    return 0;
}
