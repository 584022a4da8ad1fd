// Weight gradient of a convolution on tcgen05 (training):
//
//   dW[co][tap][ci] = sum over output pixels p of  dY[p, co] * X[p*stride + tap - pad, ci]
//
// GEMM view: M = Cout (tiles of 128), N = Cin (tiles of BN <= 128), K = output pixels, one (tap, M tile,
// N tile, K split) per work item. Both operands are NHWC activation tensors, i.e. the contraction index
// (pixels) is the SLOW axis of each tile: they are fed to the tensor cores as MN-major operands
// (instruction-descriptor bits 15/16). A TMA box of (64 channels x 128 pixels) lands in shared memory as
// 128 rows of 128 B with SWIZZLE_128B, which IS the canonical MN-major layout: 8 pixel rows x 64 channels per
// swizzle atom, SBO = 1024 B between 8-pixel groups, LBO = 16 KB between 64-channel atoms. One k-iteration =
// 128 pixels = 8 UMMA K steps (16 pixels = 2 KB of rows each). The X box is the dY box shifted by the tap
// (zero padding = TMA out-of-bounds fill; stride-2 convs read the parity planes of X exactly as in
// vu_conv_igemm.cu). Partial sums of the K splits are combined with fp32 atomics into a zeroed dW.
#include "vu_conv_common.cuh"

namespace vu {

constexpr int kWgThreads = 192;
constexpr int kWgPix = 128;                       // pixels per k-iteration
constexpr int kWgAtomBytes = kWgPix * 128;        // one 64-channel atom: 16 KB
constexpr int kWgABytes = 2 * kWgAtomBytes;       // M = 128 output channels

struct WgradParams {
    int wt, ht, nt, tiles_w, tiles_h, tiles_n, num_pix_tiles;
    int num_taps, num_m_tiles, num_n_tiles, ksplit, tiles_per_split;
    int cout, cin;
    int8_t tap_dh[kMaxTapsCommon], tap_dw[kMaxTapsCommon], tap_map[kMaxTapsCommon];
    float* dw;                                    // fp32, accumulated atomically: dw[co*co_stride + tap*tap_stride + ci*ci_stride]
    int co_stride, tap_stride, ci_stride;         // packed [Cout][taps][Cin] or OIHW [Cout][Cin][taps]
    int dtype;
};

template <int BN> struct WgCfg {
    static constexpr int kBBytes = (BN / 64) * kWgAtomBytes;
    static constexpr int kStageBytes = kWgABytes + kBBytes;
    static constexpr int kStages = (BN == 256) ? 2 : ((BN == 128) ? 3 : 4);     // BN = 256: 96 KB per stage
    static constexpr int kTmemCols = BN < 32 ? 32 : BN;
    static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + 256;
};

// MN-major, 128-byte-swizzled operand: start>>4 [0,14), LBO>>4 [16,30) = stride between 64-element MN atoms,
// SBO>>4 [32,46) = stride between 8-row K groups, version 1 [46,48), layout 2 (SWIZZLE_128B) [61,64).
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFF);
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= static_cast<uint64_t>(1024 >> 4) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}

template <int BN>
__global__ void __launch_bounds__(kWgThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap map_dy, const __grid_constant__ CUtensorMap map_x0,
                  const __grid_constant__ CUtensorMap map_x1, const __grid_constant__ CUtensorMap map_x2,
                  const __grid_constant__ CUtensorMap map_x3, const WgradParams p) {
    using Cfg = WgCfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
    const uint32_t bar_base = smem_base + Cfg::kStages * Cfg::kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (Cfg::kStages + s); };
    const uint32_t done_bar = bar_base + 8u * (2 * Cfg::kStages);
    const uint32_t tmem_slot = bar_base + 8u * (2 * Cfg::kStages + 1);
    volatile uint32_t* tmem_slot_ptr =
        reinterpret_cast<volatile uint32_t*>(smem_al + Cfg::kStages * Cfg::kStageBytes + 8 * (2 * Cfg::kStages + 1));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_dy);
        tma_prefetch_desc(&map_x0);
        for (int s = 0; s < Cfg::kStages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        mbar_init(done_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc<Cfg::kTmemCols>(tmem_slot);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    // work item = blockIdx.x -> (ksplit, n_tile, m_tile, tap), tap fastest (neighbouring CTAs share pixel tiles in L2)
    int item = blockIdx.x;
    const int tap = item % p.num_taps; item /= p.num_taps;
    const int m_tile = item % p.num_m_tiles; item /= p.num_m_tiles;
    const int n_tile = item % p.num_n_tiles; item /= p.num_n_tiles;
    const int ks = item;
    const int t_begin = ks * p.tiles_per_split;
    const int t_end = min(p.num_pix_tiles, t_begin + p.tiles_per_split);
    const int n_iters = max(0, t_end - t_begin);

    if (warp == 0) {
        if (elect_one()) {
            const CUtensorMap* maps[4] = {&map_x0, &map_x1, &map_x2, &map_x3};
            const CUtensorMap* mx = maps[p.tap_map[tap]];
            for (int i = 0; i < n_iters; ++i) {
                const int t = t_begin + i;
                const int wb = t % p.tiles_w;
                const int hb = (t / p.tiles_w) % p.tiles_h;
                const int nb = t / (p.tiles_w * p.tiles_h);
                const int w0 = wb * p.wt, h0 = hb * p.ht, n0 = nb * p.nt;
                const int s = i % Cfg::kStages;
                mbar_wait(empty_bar(s), ((i / Cfg::kStages) & 1) ^ 1);
                const uint32_t sa = smem_base + s * Cfg::kStageBytes;
                mbar_arrive_expect_tx(full_bar(s), Cfg::kStageBytes);
                tma_load_4d(sa, &map_dy, full_bar(s), m_tile * 128, w0, h0, n0);
                tma_load_4d(sa + kWgAtomBytes, &map_dy, full_bar(s), m_tile * 128 + 64, w0, h0, n0);
#pragma unroll
                for (int a = 0; a < BN / 64; ++a)
                    tma_load_4d(sa + kWgABytes + a * kWgAtomBytes, mx, full_bar(s), n_tile * BN + a * 64,
                                w0 + p.tap_dw[tap], h0 + p.tap_dh[tap], n0);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // bits 15/16: A and B are MN-major
        const uint32_t idesc = umma_idesc_f16(128, BN, p.dtype == 0 ? 1u : 0u) | (1u << 15) | (1u << 16);
        for (int i = 0; i < n_iters; ++i) {
            const int s = i % Cfg::kStages;
            mbar_wait(full_bar(s), (i / Cfg::kStages) & 1);
            tc_fence_after();
            const uint32_t sa = smem_base + s * Cfg::kStageBytes;
            const uint64_t adesc = umma_desc_mn_sw128(sa, kWgAtomBytes);
            const uint64_t bdesc = umma_desc_mn_sw128(sa + kWgABytes, kWgAtomBytes);
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < kWgPix / 16; ++k)       // 16 pixel rows = 2048 B per K step (>>4 = 128)
                    umma_f16_ss(tmem_base, adesc + 128u * k, bdesc + 128u * k, idesc, (i | k) != 0);
                umma_commit(empty_bar(s));
                if (i == n_iters - 1) umma_commit(done_bar);
            }
            __syncwarp();
        }
    } else if (n_iters > 0) {
        // epilogue: lane of TMEM = output channel, columns = input channels
        const int q = warp & 3;
        const int co = m_tile * 128 + q * 32 + lane;
        mbar_wait(done_bar, 0);
        tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
            uint32_t r[32];
            tmem_ld_32x32b_x32(taddr + c * 32, r);
            tmem_ld_wait();
            const int ci0 = n_tile * BN + c * 32;
            if (co < p.cout) {
                float* dst = p.dw + static_cast<size_t>(co) * p.co_stride + tap * p.tap_stride + static_cast<size_t>(ci0) * p.ci_stride;
                if (p.ci_stride == 1) {
                    // packed layout: this thread's 32 input channels are 128 contiguous bytes -> eight 16-byte vector reductions
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        if (ci0 + j < p.cin)
                            red_add_v4_f32(dst + j, __uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                                           __uint_as_float(r[j + 3]));
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (ci0 + j < p.cin) atomicAdd(dst + j * p.ci_stride, __uint_as_float(r[j]));
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<Cfg::kTmemCols>(tmem_base);
    }
}


// ------------------------------------------------------------------------------------------------------------------
// 3x3 stride-1 weight gradient, one filter ROW (3 taps) per work item. The per-tap kernel above re-reads the dY tile and the
// X tile for every tap and is bound by L2->SM traffic (ncu: 9.5 TB/s, tensor pipe < 25 %). Here a k-iteration loads the dY
// tile once and the X tile once with a one-pixel halo on both sides of every tile row (box (64 ch, wt+2, ht, nt) at
// w0-1, h0+r-1), and issues the MMAs of the three taps s = 0,1,2 from it: tap s reads the same shared-memory tile with
// the operand descriptor advanced by s pixel rows (128 B each; the 128-byte swizzle is a function of the absolute
// shared-memory address, the property vu_conv_halo.cu relies on). Three fp32 accumulators of BN columns live in TMEM.
// Traffic per 3 taps: 32 KB (dY) + ~36 KB (X) instead of 3 x 64 KB.
//
// Cout <= 64 (`stack`): an M = 64 MMA costs the same cycles as M = 128, and a zero-filled upper half of the dY tile wasted half
// of every MMA on the 64-channel layers (layer1, the last decoder block: 1.3 ms of weight gradients per step). Instead the
// second 64 rows of the A tile are the SAME 64 channels of dY one image row UP (box at h0 - 1): with X at row h + r - 1,
//   rows  0..63  accumulate  sum dY[h]   * X[h + r - 1]  = filter row r,
//   rows 64..127 accumulate  sum dY[h-1] * X[h + r - 1]  = sum dY[h'] * X[h' + r] = filter row r + 1,
// so the work items are r in {0, 2} (two instead of three; the upper half of r = 2 is discarded) and the tile grid covers one
// extra image row (h = H: dY out of bounds = TMA zero fill for the lower half, dY[H-1] for the upper half).
struct WgradRow3Params {
    int wt, ht, nt, tiles_w, tiles_h, tiles_n, num_pix_tiles;
    int num_m_tiles, num_n_tiles, ksplit, tiles_per_split;
    int cout, cin;
    int b_atom_bytes;        // (wt+2)*ht*nt*128 rounded up to 1024
    int b_box_bytes;         // (wt+2)*ht*nt*128
    int stage_bytes, stages;
    float* dw;               // packed [Cout][9][Cin]
    int dtype;
    int stack;               // Cout <= 64: two filter rows per work item (see above)
};

template <int BN>
__global__ void __launch_bounds__(kWgThreads, 1)
conv_wgrad_row3_kernel(const __grid_constant__ CUtensorMap map_dy, const __grid_constant__ CUtensorMap map_xh,
                       const WgradRow3Params p) {
    constexpr int kTmemCols = 512;                 // 3 accumulators x BN (<= 128) columns
    constexpr int kMaxStages = 4;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
    const int stages = p.stages;
    const uint32_t bar_base = smem_base + stages * p.stage_bytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (kMaxStages + s); };
    const uint32_t done_bar = bar_base + 8u * (2 * kMaxStages);
    const uint32_t tmem_slot = bar_base + 8u * (2 * kMaxStages + 1);
    volatile uint32_t* tmem_slot_ptr =
        reinterpret_cast<volatile uint32_t*>(smem_al + stages * p.stage_bytes + 8 * (2 * kMaxStages + 1));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_dy);
        tma_prefetch_desc(&map_xh);
        for (int s = 0; s < stages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        mbar_init(done_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc<kTmemCols>(tmem_slot);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    // work item = blockIdx.x -> (ksplit, n_tile, m_tile, r), r fastest (neighbouring CTAs share pixel tiles in L2)
    int item = blockIdx.x;
    int r;
    if (p.stack) { r = (item & 1) * 2; item >>= 1; }
    else { r = item % 3; item /= 3; }
    const int m_tile = item % p.num_m_tiles; item /= p.num_m_tiles;
    const int n_tile = item % p.num_n_tiles; item /= p.num_n_tiles;
    const int ks = item;
    const int t_begin = ks * p.tiles_per_split;
    const int t_end = min(p.num_pix_tiles, t_begin + p.tiles_per_split);
    const int n_iters = max(0, t_end - t_begin);
    const int pitch = p.wt + 2;                    // pixels per tile row of the halo'd X tile

    if (warp == 0) {
        if (elect_one()) {
            for (int i = 0; i < n_iters; ++i) {
                const int t = t_begin + i;
                const int wb = t % p.tiles_w;
                const int hb = (t / p.tiles_w) % p.tiles_h;
                const int nb = t / (p.tiles_w * p.tiles_h);
                const int w0 = wb * p.wt, h0 = hb * p.ht, n0 = nb * p.nt;
                const int s = i % stages;
                mbar_wait(empty_bar(s), ((i / stages) & 1) ^ 1);
                const uint32_t sa = smem_base + s * p.stage_bytes;
                mbar_arrive_expect_tx(full_bar(s), kWgABytes + (BN / 64) * p.b_box_bytes);
                tma_load_4d(sa, &map_dy, full_bar(s), m_tile * 128, w0, h0, n0);
                if (p.stack) tma_load_4d(sa + kWgAtomBytes, &map_dy, full_bar(s), 0, w0, h0 - 1, n0);
                else tma_load_4d(sa + kWgAtomBytes, &map_dy, full_bar(s), m_tile * 128 + 64, w0, h0, n0);
#pragma unroll
                for (int a = 0; a < BN / 64; ++a)
                    tma_load_4d(sa + kWgABytes + a * p.b_atom_bytes, &map_xh, full_bar(s), n_tile * BN + a * 64, w0 - 1,
                                h0 + r - 1, n0);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        const uint32_t idesc = umma_idesc_f16(128, BN, p.dtype == 0 ? 1u : 0u) | (1u << 15) | (1u << 16);
        for (int i = 0; i < n_iters; ++i) {
            const int s = i % stages;
            mbar_wait(full_bar(s), (i / stages) & 1);
            tc_fence_after();
            const uint32_t sa = smem_base + s * p.stage_bytes;
            const uint64_t adesc = umma_desc_mn_sw128(sa, kWgAtomBytes);
            const uint64_t bdesc = umma_desc_mn_sw128(sa + kWgABytes, p.b_atom_bytes);
            if (elect_one()) {
#pragma unroll 1
                for (int k = 0; k < kWgPix / 16; ++k) {     // 16 tile pixels (inside one tile row: wt % 16 == 0)
                    const int lin = k * 16;
                    const uint32_t brow = (lin / p.wt) * pitch + (lin % p.wt);     // halo'd row index of tap s = 0
#pragma unroll
                    for (int sx = 0; sx < 3; ++sx)
                        umma_f16_ss(tmem_base + sx * BN, adesc + 128u * k, bdesc + 8u * (brow + sx), idesc, (i | k) != 0);
                }
                umma_commit(empty_bar(s));
                if (i == n_iters - 1) umma_commit(done_bar);
            }
            __syncwarp();
        }
    } else if (n_iters > 0) {
        const int q = warp & 3;
        const int co = p.stack ? (q & 1) * 32 + lane : m_tile * 128 + q * 32 + lane;
        const int tr = p.stack ? r + (q >> 1) : r;           // filter row this TMEM lane quarter holds
        mbar_wait(done_bar, 0);
        tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
        for (int sx = 0; sx < 3 && tr < 3; ++sx) {
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                uint32_t v[32];
                tmem_ld_32x32b_x32(taddr + sx * BN + c * 32, v);
                tmem_ld_wait();
                const int ci0 = n_tile * BN + c * 32;
                if (co < p.cout) {
                    float* dst = p.dw + (static_cast<size_t>(co) * 9 + tr * 3 + sx) * p.cin + ci0;
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        if (ci0 + j < p.cin)
                            red_add_v4_f32(dst + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                           __uint_as_float(v[j + 3]));
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
}

}  // namespace vu

using namespace vu;

static void wg_choose_tile(int n, int h, int w, int* wt_o, int* ht_o, int* nt_o) {
    double best = -1.0;
    for (int wt = 128; wt >= 1; wt >>= 1)
        for (int ht = 128 / wt; ht >= 1; ht >>= 1) {
            const int nt = 128 / (wt * ht);
            const double cover = (double)((w + wt - 1) / wt * wt) * ((h + ht - 1) / ht * ht) * ((n + nt - 1) / nt * nt);
            const double eff = (double)w * h * n / cover;
            if (eff > best + 1e-9) { best = eff; *wt_o = wt; *ht_o = ht; *nt_o = nt; }
        }
}

template <int BN>
static int launch_wgrad(const CUtensorMap& dym, const CUtensorMap* xm, const WgradParams& wp, int grid, cudaStream_t stream) {
    using Cfg = WgCfg<BN>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv_wgrad_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "cudaFuncSetAttribute(wgrad): %s", cudaGetErrorString(e));
        attr_set = true;
    }
    conv_wgrad_kernel<BN><<<grid, kWgThreads, Cfg::kSmemBytes, stream>>>(dym, xm[0], xm[1], xm[2], xm[3], wp);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "conv_wgrad launch: %s", cudaGetErrorString(e));
    return 0;
}

// K split of a weight-gradient launch. Default: a cost model - time ~ waves x (k-iterations per CTA + c), c = the prologue +
// epilogue (TMEM allocation, pipeline fill, the 128 x BN fp32 reductions) in units of one k-iteration (VU_WGRAD_OVERHEAD,
// default 12; 3..24 measure within 0.05 ms of each other, profiles/r2zq_ksplit.log) - minimised over the split count. The first policy, "fill 1.5 waves" (VU_WGRAD_HALFWAVES=n selects it for A/B
// runs), overshot whole waves: 9 base items -> 25 splits = 225 CTAs = 1.52 waves of 20 k-iterations, where 16 splits = 144
// CTAs = ONE wave of 32 k-iterations does the same work in ~25 % less time with a third less reduction traffic. Whole training step 12.43 -> 11.89 ms.
static int wg_choose_ksplit(int base_items, int num_pix_tiles, int* tiles_per_split_out) {
    static const int halfwaves = getenv("VU_WGRAD_HALFWAVES") ? atoi(getenv("VU_WGRAD_HALFWAVES")) : 0;
    static const double c_over = getenv("VU_WGRAD_OVERHEAD") ? atof(getenv("VU_WGRAD_OVERHEAD")) : 12.0;
    int ksplit = 1;
    if (halfwaves > 0) {
        ksplit = (halfwaves * num_sms() / 2 + base_items - 1) / base_items;
        if (ksplit > (num_pix_tiles + 3) / 4) ksplit = (num_pix_tiles + 3) / 4;
        if (ksplit < 1) ksplit = 1;
    } else {
        double best = 1e30;
        const int sms = num_sms();
        const int ks_max = num_pix_tiles < 256 ? num_pix_tiles : 256;
        for (int ks = 1; ks <= ks_max; ++ks) {
            const int tps = (num_pix_tiles + ks - 1) / ks;
            const int ks_eff = (num_pix_tiles + tps - 1) / tps;
            const long ctas = (long)base_items * ks_eff;
            const long waves = (ctas + sms - 1) / sms;
            const double cost = (double)waves * (tps + c_over);
            if (cost < best - 1e-9) { best = cost; ksplit = ks_eff; }
        }
    }
    const int tps = (num_pix_tiles + ksplit - 1) / ksplit;
    *tiles_per_split_out = tps;
    return (num_pix_tiles + tps - 1) / tps;
}

// tile (wt, ht, nt) with wt a multiple of 16 (a 16-pixel MMA K step must stay inside one tile row), product 128
static bool wg_choose_tile_row3(int n, int h, int w, int* wt_o, int* ht_o, int* nt_o) {
    double best = -1.0;
    for (int wt = 128; wt >= 16; wt >>= 1)
        for (int ht = 128 / wt; ht >= 1; ht >>= 1) {
            const int nt = 128 / (wt * ht);
            const double cover = (double)((w + wt - 1) / wt * wt) * ((h + ht - 1) / ht * ht) * ((n + nt - 1) / nt * nt);
            const double eff = (double)w * h * n / cover;
            if (eff > best + 1e-9) { best = eff; *wt_o = wt; *ht_o = ht; *nt_o = nt; }
        }
    return best >= 0.75;        // narrow / ragged maps stay on the per-tap kernel
}

template <int BN>
static int launch_wgrad_row3(const CUtensorMap& dym, const CUtensorMap& xhm, const WgradRow3Params& wp, int grid, size_t smem,
                             cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv_wgrad_row3_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "cudaFuncSetAttribute(wgrad row3): %s", cudaGetErrorString(e));
        attr_set = true;
    }
    conv_wgrad_row3_kernel<BN><<<grid, kWgThreads, smem, stream>>>(dym, xhm, wp);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "conv_wgrad_row3 launch: %s", cudaGetErrorString(e));
    return 0;
}

// returns 1 when it took the launch, 0 when the shape is not its case, < 0 on error
static int wgrad_row3_try(const vu_conv_args* a, const void* dy, int dy_ctot, int dy_coff, float* dw, cudaStream_t stream) {
    if (!(a->KH == 3 && a->KW == 3 && a->stride == 1 && a->pad == 1)) return 0;
    if (getenv("VU_WGRAD_PER_TAP")) return 0;
    WgradRow3Params wp{};
    if (!wg_choose_tile_row3(a->N, a->H, a->W, &wp.wt, &wp.ht, &wp.nt)) return 0;
    const int Ho = a->H, Wo = a->W;
    static const bool no_stack = getenv("VU_WGRAD_NO_STACK") != nullptr;
    wp.stack = (!no_stack && a->Cout <= 64 && wp.ht * 8 <= Ho) ? 1 : 0;
    wp.tiles_w = (Wo + wp.wt - 1) / wp.wt; wp.tiles_h = (Ho + wp.stack + wp.ht - 1) / wp.ht; wp.tiles_n = (a->N + wp.nt - 1) / wp.nt;
    wp.num_pix_tiles = wp.tiles_w * wp.tiles_h * wp.tiles_n;
    // measured on B200 (scripts/prof_wgrad.py): the row kernel wins from 2048 pixel tiles up (+30..60 %), the per-tap kernel
    // below 512 (three times more work items keep the K splits long enough to amortise prologue + reductions)
    static const int row3_min = getenv("VU_WGRAD_ROW3_MIN") ? atoi(getenv("VU_WGRAD_ROW3_MIN")) : 512;   // 1024 -> 512 with the cost-model K split: step 11.84 -> 11.77 ms (profiles/r2zs_row3min.log)
    if (wp.num_pix_tiles < row3_min && !getenv("VU_WGRAD_ROW3")) return 0;
    wp.cout = a->Cout; wp.cin = a->Cin; wp.dw = dw; wp.dtype = a->dtype;
    const int bn = (a->Cin > 64) ? 128 : 64;
    wp.num_m_tiles = (a->Cout + 127) / 128;
    wp.num_n_tiles = (a->Cin + bn - 1) / bn;
    wp.b_box_bytes = (wp.wt + 2) * wp.ht * wp.nt * 128;
    wp.b_atom_bytes = (wp.b_box_bytes + 1023) / 1024 * 1024;
    wp.stage_bytes = kWgABytes + (bn / 64) * wp.b_atom_bytes;
    const int smem_budget = 227 * 1024 - 1024 /*align slack*/ - 256 /*barriers*/;
    wp.stages = smem_budget / wp.stage_bytes;
    if (wp.stages > 4) wp.stages = 4;
    if (wp.stages < 2) return 0;
    const int base_items = (wp.stack ? 2 : 3) * wp.num_m_tiles * wp.num_n_tiles;
    wp.ksplit = wg_choose_ksplit(base_items, wp.num_pix_tiles, &wp.tiles_per_split);

    CUtensorMap dym, xhm;
    {
        const uint32_t box[4] = {64u, (uint32_t)wp.wt, (uint32_t)wp.ht, (uint32_t)wp.nt};
        const uint16_t* db = static_cast<const uint16_t*>(dy) + dy_coff;
        const uint64_t pix = (uint64_t)dy_ctot * 2;
        const uint64_t dims[4] = {(uint64_t)a->Cout, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)a->N};
        const uint64_t str[3] = {pix, pix * Wo, pix * Wo * Ho};
        int rc = encode_map(&dym, a->dtype, 4, db, dims, str, box);
        if (rc) return rc;
    }
    {
        const uint32_t box[4] = {64u, (uint32_t)(wp.wt + 2), (uint32_t)wp.ht, (uint32_t)wp.nt};
        const uint16_t* xb = static_cast<const uint16_t*>(a->x) + a->x_coff;
        const uint64_t pix = (uint64_t)a->x_ctot * 2;
        const uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->N};
        const uint64_t str[3] = {pix, pix * a->W, pix * a->W * a->H};
        int rc = encode_map(&xhm, a->dtype, 4, xb, dims, str, box);
        if (rc) return rc;
    }
    const int grid = base_items * wp.ksplit;
    const size_t smem = 1024 + (size_t)wp.stages * wp.stage_bytes + 256;
    int rc = (bn == 128) ? launch_wgrad_row3<128>(dym, xhm, wp, grid, smem, stream)
                         : launch_wgrad_row3<64>(dym, xhm, wp, grid, smem, stream);
    return rc ? rc : 1;
}

static int wgrad_impl(const vu_conv_args* a, const void* dy, int dy_ctot, int dy_coff, float* dw, bool oihw,
                      vu_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!a || !a->x || !dy || !dw) return set_error(VU_ERR_INVALID, "wgrad: null pointer");
    const int KH = a->KH, KW = a->KW, S = a->stride, P = a->pad;
    const bool ok_geom = (KH == KW) && ((KH == 3 && P == 1) || (KH == 1 && P == 0)) && (S == 1 || S == 2);
    if (!ok_geom) return set_error(VU_ERR_UNSUPPORTED, "wgrad: only 3x3 p1 / 1x1 p0, stride 1|2 (got k%d s%d p%d)", KH, S, P);
    if (a->Cin % 8 || a->x_ctot % 8 || a->x_coff % 8 || dy_ctot % 8 || dy_coff % 8 || a->Cout % 8)
        return set_error(VU_ERR_ALIGN, "wgrad: channel counts/offsets must be multiples of 8");
    if ((reinterpret_cast<uintptr_t>(a->x) | reinterpret_cast<uintptr_t>(dy)) & 15)
        return set_error(VU_ERR_ALIGN, "wgrad: pointers must be 16-byte aligned");
    const int Ho = (a->H + 2 * P - KH) / S + 1, Wo = (a->W + 2 * P - KW) / S + 1;
    if (!oihw) {
        const int rc = wgrad_row3_try(a, dy, dy_ctot, dy_coff, dw, stream);
        if (rc != 0) return rc < 0 ? rc : 0;
    }

    WgradParams wp{};
    wg_choose_tile(a->N, Ho, Wo, &wp.wt, &wp.ht, &wp.nt);
    wp.tiles_w = (Wo + wp.wt - 1) / wp.wt; wp.tiles_h = (Ho + wp.ht - 1) / wp.ht; wp.tiles_n = (a->N + wp.nt - 1) / wp.nt;
    wp.num_pix_tiles = wp.tiles_w * wp.tiles_h * wp.tiles_n;
    wp.num_taps = KH * KW;
    wp.cout = a->Cout; wp.cin = a->Cin;
    wp.dw = dw; wp.dtype = a->dtype;
    if (oihw) { wp.co_stride = a->Cin * KH * KW; wp.tap_stride = 1; wp.ci_stride = KH * KW; }
    else { wp.co_stride = KH * KW * a->Cin; wp.tap_stride = a->Cin; wp.ci_stride = 1; }
    // N tile: 256 input channels per work item (VU_WGRAD_BN256=1) was built against the L2->SM bound of this kernel (ncu r1c:
    // 8.2 TB/s): a k-iteration then moves 32 KB (dY) + 64 KB (X) for twice the MMA work of 32 + 32 KB. In the step it measured
    // no gain (in-situ CUPTI sums 2.12 vs 2.01 ms per step over the 37 per-tap launches: only two 96 KB stages fit, and at
    // B = 16 these launches are dominated by their split-K reductions, not by operand traffic), so it stays off by default.
    const int bn256 = getenv("VU_WGRAD_BN256") ? atoi(getenv("VU_WGRAD_BN256")) : 0;
    const int bn = (a->Cin > 64) ? ((bn256 && a->Cin % 256 == 0) ? 256 : 128) : 64;
    wp.num_m_tiles = (a->Cout + 127) / 128;
    wp.num_n_tiles = (a->Cin + bn - 1) / bn;
    const int base_items = wp.num_taps * wp.num_m_tiles * wp.num_n_tiles;
    // K splits: cost model (wg_choose_ksplit)
    wp.ksplit = wg_choose_ksplit(base_items, wp.num_pix_tiles, &wp.tiles_per_split);

    const uint32_t box[4] = {64u, (uint32_t)wp.wt, (uint32_t)wp.ht, (uint32_t)wp.nt};
    CUtensorMap dym, xm[4];
    {
        const uint16_t* db = static_cast<const uint16_t*>(dy) + dy_coff;
        const uint64_t pix = (uint64_t)dy_ctot * 2;
        const uint64_t dims[4] = {(uint64_t)a->Cout, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)a->N};
        const uint64_t str[3] = {pix, pix * Wo, pix * Wo * Ho};
        int rc = encode_map(&dym, a->dtype, 4, db, dims, str, box);
        if (rc) return rc;
    }
    const uint16_t* xb = static_cast<const uint16_t*>(a->x) + a->x_coff;
    const uint64_t pix = (uint64_t)a->x_ctot * 2;
    if (S == 1) {
        const uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->N};
        const uint64_t str[3] = {pix, pix * a->W, pix * a->W * a->H};
        int rc = encode_map(&xm[0], a->dtype, 4, xb, dims, str, box);
        if (rc) return rc;
        xm[1] = xm[2] = xm[3] = xm[0];
        int t = 0;
        for (int r = 0; r < KH; ++r)
            for (int s = 0; s < KW; ++s, ++t) {
                wp.tap_dh[t] = (int8_t)(r - P); wp.tap_dw[t] = (int8_t)(s - P); wp.tap_map[t] = 0;
            }
    } else {
        for (int pr = 0; pr < 2; ++pr)
            for (int pc = 0; pc < 2; ++pc) {
                const int hp = (a->H - pr + 1) / 2, wpn = (a->W - pc + 1) / 2;
                const int id = pr * 2 + pc;
                if (hp <= 0 || wpn <= 0) { xm[id] = xm[0]; continue; }
                const uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)wpn, (uint64_t)hp, (uint64_t)a->N};
                const uint64_t str[3] = {pix * 2, pix * a->W * 2, pix * a->W * a->H};
                int rc = encode_map(&xm[id], a->dtype, 4, xb + ((size_t)pr * a->W + pc) * a->x_ctot, dims, str, box);
                if (rc) return rc;
            }
        int t = 0;
        for (int r = 0; r < KH; ++r)
            for (int s = 0; s < KW; ++s, ++t) {
                const int ir = r - P, ic = s - P;
                const int pr = ((ir % 2) + 2) % 2, pc = ((ic % 2) + 2) % 2;
                wp.tap_dh[t] = (int8_t)((ir - pr) / 2); wp.tap_dw[t] = (int8_t)((ic - pc) / 2);
                wp.tap_map[t] = (int8_t)(pr * 2 + pc);
            }
    }
    const int grid = base_items * wp.ksplit;
    if (bn == 256) return launch_wgrad<256>(dym, xm, wp, grid, stream);
    if (bn == 128) return launch_wgrad<128>(dym, xm, wp, grid, stream);
    return launch_wgrad<64>(dym, xm, wp, grid, stream);
}

extern "C" int vu_conv2d_wgrad(const vu_conv_args* a, const void* dy, int dy_ctot, int dy_coff, float* dw, vu_stream_t stream) {
    return wgrad_impl(a, dy, dy_ctot, dy_coff, dw, false, stream);
}

extern "C" int vu_conv2d_wgrad_oihw(const vu_conv_args* a, const void* dy, int dy_ctot, int dy_coff, float* dw, vu_stream_t stream) {
    return wgrad_impl(a, dy, dy_ctot, dy_coff, dw, true, stream);
}
