// Implicit-GEMM convolution for sm_100a: tcgen05.mma (UMMA 128 x BN x 16, bf16/fp16 -> fp32 in
// TMEM), operands staged by TMA with 128-byte swizzle, persistent warp-specialised CTAs.
//
//   GEMM view:  M = output pixels (tiles of 128 = Nt x Ht x Wt pixels of the NHWC map),
//               N = Cout (tiles of BN), K = taps x Cin (blocks of 64 channels per tap).
//   A operand:  for tap (r,s) the 128 x 64 tile is ONE 4-D TMA box (64ch, Wt, Ht, Nt) of the NHWC
//               activation tensor shifted by (r-pad, s-pad); zero padding = TMA out-of-bounds fill.
//               Stride-2 convs read the four (row,col)-parity planes of the input through four
//               strided tensor maps, so every tap is again a unit-stride box.
//   B operand:  packed weights [Cout][tap][Cin] through a 3-D map, box (64ch, 1, BN).
//   Both land in smem as rows of 128 B in the canonical K-major SWIZZLE_128B layout.
//   Roles:      warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (one lane),
//               warps 2..5 = epilogue (TMEM -> regs -> scale/bias/residual/ReLU -> 16-bit NHWC slice).
//   Pipelines:  smem full/empty ring (kStages), TMEM full/empty (2 accumulators) so the epilogue of
//               tile i overlaps the MMAs of tile i+1.
#include "vu_conv_common.cuh"

namespace vu {

constexpr int kTileM = 128;
constexpr int kBlockK = 64;                       // channels per K block (128 B rows)
constexpr int kABytes = kTileM * kBlockK * 2;     // 16 KB
constexpr int kMaxTaps = 9;
constexpr int kMaxCout = 1024;
// Warps: 0 = TMA producer, 1 = TMEM allocator + MMA issuer, 2..9 = epilogue. EIGHT epilogue warps (two per TMEM lane quarter,
// taking the even / odd 32-column chunks of a tile): with K = 9 * 64..256 the MMAs of a 128 x 128 tile take ~4.6k clk, and one
// warp per scheduler needed ~12k clk to drain it (ncu r2b: the MMA warp sat on the accumulator-empty barrier, tensor pipe 24 %).
constexpr int kEpiWarps = 8;
constexpr int kThreads = 64 + 32 * kEpiWarps;
constexpr int kStatMaxCout = kEpiStatMaxCout;     // BatchNorm statistics in the epilogue: Cout <= 512
constexpr int kSmemLimit = 232448;

struct ConvKParams {
    int n_img, ho, wo;             // output-resolution grid the M tiles cover
    int wt, ht, nt;                // tile = nt x ht x wt pixels (product 128)
    int tiles_w, tiles_h, tiles_n;
    int num_m_tiles, num_n_tiles;
    int cout;
    int num_taps, kblocks, last_ksteps;
    int8_t tap_dh[kMaxTaps], tap_dw[kMaxTaps], tap_map[kMaxTaps], tap_widx[kMaxTaps];
    const float* scale; const float* bias;
    const uint16_t* residual; int res_ctot, res_coff;
    uint16_t* y; int y_ctot, y_coff;
    int out_h, out_w, out_up, out_dh, out_dw;
    int relu, dtype;
    const float* pos_bias;         // optional [n_img][9][cout] pre-scale bias by border class (3x3 s1 p1)
    double* stat_sums;             // optional [2*cout] fp64: sum / sum of squares of the stored outputs
};

// STATS: the kernel flavour that accumulates BatchNorm statistics (training): it carries the per-CTA partial sums and the
// per-warp transpose scratch in shared memory and therefore one or two pipeline stages fewer.
template <int BN, bool STATS> struct ConvCfg {
    static constexpr int kBBytes = BN * kBlockK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kBarBytes = 1024;  // barriers + tmem ptr
    static constexpr int kTailBytes = kBarBytes + 2 * kMaxCout * 4 + (STATS ? kEpiStatBytes + kEpiWarps * kEpiScratchBytes : 0);
    static constexpr int kMaxStages = (kSmemLimit - 1024 - kTailBytes) / kStageBytes;
    static constexpr int kStages = kMaxStages > 8 ? 8 : kMaxStages;
    static constexpr int kTmemCols = (2 * BN < 32) ? 32 : 2 * BN;
    static constexpr int kSmemBytes = 1024 /*align slack*/ + kStages * kStageBytes + kTailBytes;
    static_assert(kStages >= 3, "pipeline depth");
};

// Epilogue of one 128-pixel x BN tile: TMEM -> registers -> (pos_bias, scale/bias, residual, ReLU) -> 16-bit NHWC store,
// plus the optional BatchNorm statistics of the stored values. Called by the eight epilogue warps: q = TMEM lane quarter,
// `half` = which 32-column chunks (even / odd) this warp drains; `taddr` addresses this warp's lanes of the accumulator.
// Every optional piece sits behind a warp-uniform BRANCH (not per-element predication): a raw training conv or a data
// gradient executes only TMEM load, pack and store.
template <int BN, bool STATS>
__device__ __forceinline__ void conv_epilogue_tile(const ConvKParams& p, uint32_t taddr, int n_tile, int m_tile, int q, int half,
                                                   int lane, const float* s_scale, const float* s_bias, float* s_stat,
                                                   uint8_t* scratch) {
    const int m = q * 32 + lane;              // row of the tile = pixel
    const int w_l = m % p.wt;
    const int h_l = (m / p.wt) % p.ht;
    const int n_l = m / (p.wt * p.ht);
    const int wb = m_tile % p.tiles_w;
    const int hb = (m_tile / p.tiles_w) % p.tiles_h;
    const int nb = m_tile / (p.tiles_w * p.tiles_h);
    const int wo = wb * p.wt + w_l, ho = hb * p.ht + h_l, ni = nb * p.nt + n_l;
    const bool valid = (wo < p.wo) && (ho < p.ho) && (ni < p.n_img);
    const size_t opix = (static_cast<size_t>(ni) * p.out_h + (ho * p.out_up + p.out_dh)) * p.out_w +
                        (wo * p.out_up + p.out_dw);
    const size_t rpix = (static_cast<size_t>(ni) * p.ho + ho) * p.wo + wo;
    const bool affine = p.scale != nullptr || p.bias != nullptr;
    const bool has_res = p.residual != nullptr, relu = p.relu != 0;
    const int dtype = p.dtype;
    constexpr int kChunks = BN / 32;
#pragma unroll 1
    for (int c = half; c < kChunks; c += 2) {
        const int ch0 = n_tile * BN + c * 32;
        if (ch0 >= p.cout) break;             // partial last N tile (Cout is a multiple of 32; warp-uniform)
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr + c * 32, r);
        uint4 rv[4];
        if (has_res) {
#pragma unroll
            for (int j = 0; j < 4; ++j) rv[j] = make_uint4(0, 0, 0, 0);
            if (valid) {
                const uint4* rp = reinterpret_cast<const uint4*>(p.residual + rpix * p.res_ctot + p.res_coff + ch0);
#pragma unroll
                for (int j = 0; j < 4; ++j) rv[j] = __ldg(rp + j);
            }
        }
        tmem_ld_wait();
        if (p.pos_bias != nullptr) {
            const int hc = min(ho, p.ho - 1), wc = min(wo, p.wo - 1), nc = min(ni, p.n_img - 1);
            const int cls = (hc == 0 ? 0 : (hc == p.ho - 1 ? 2 : 1)) * 3 + (wc == 0 ? 0 : (wc == p.wo - 1 ? 2 : 1));
            const float4* pp = reinterpret_cast<const float4*>(p.pos_bias + (static_cast<size_t>(nc) * 9 + cls) * p.cout + ch0);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 b = __ldg(pp + j);
                r[4 * j] = __float_as_uint(__uint_as_float(r[4 * j]) + b.x);
                r[4 * j + 1] = __float_as_uint(__uint_as_float(r[4 * j + 1]) + b.y);
                r[4 * j + 2] = __float_as_uint(__uint_as_float(r[4 * j + 2]) + b.z);
                r[4 * j + 3] = __float_as_uint(__uint_as_float(r[4 * j + 3]) + b.w);
            }
        }
        if (affine) {
            const float4* sc4 = reinterpret_cast<const float4*>(s_scale + ch0);
            const float4* sb4 = reinterpret_cast<const float4*>(s_bias + ch0);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 a = sc4[j], b = sb4[j];
                r[4 * j] = __float_as_uint(fmaf(__uint_as_float(r[4 * j]), a.x, b.x));
                r[4 * j + 1] = __float_as_uint(fmaf(__uint_as_float(r[4 * j + 1]), a.y, b.y));
                r[4 * j + 2] = __float_as_uint(fmaf(__uint_as_float(r[4 * j + 2]), a.z, b.z));
                r[4 * j + 3] = __float_as_uint(fmaf(__uint_as_float(r[4 * j + 3]), a.w, b.w));
            }
        }
        if (has_res) {
            const uint32_t* rw = reinterpret_cast<const uint32_t*>(rv);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float2 rr = unpack2(rw[j], dtype);
                r[2 * j] = __float_as_uint(__uint_as_float(r[2 * j]) + rr.x);
                r[2 * j + 1] = __float_as_uint(__uint_as_float(r[2 * j + 1]) + rr.y);
            }
        }
        if (relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(fmaxf(__uint_as_float(r[j]), 0.0f));
        }
        uint32_t outw[16];
        if (dtype == 0) {
#pragma unroll
            for (int j = 0; j < 16; ++j) outw[j] = pack2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]), 0);
        } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) outw[j] = pack2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]), 1);
        }
        if (valid) {
            uint4* op = reinterpret_cast<uint4*>(p.y + opix * p.y_ctot + p.y_coff + ch0);
#pragma unroll
            for (int j = 0; j < 4; ++j)
                op[j] = make_uint4(outw[4 * j], outw[4 * j + 1], outw[4 * j + 2], outw[4 * j + 3]);
        }
        if constexpr (STATS) {
            if (p.stat_sums != nullptr)        // statistics of the values as stored (16-bit rounded); warp-uniform branch
                epi_stats_chunk(outw, valid, lane, dtype, scratch, s_stat + (q * 2) * kStatMaxCout + ch0,
                                s_stat + (q * 2 + 1) * kStatMaxCout + ch0);
        }
    }
}

template <int BN, bool STATS>
__global__ void __launch_bounds__(kThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap map_a0, const __grid_constant__ CUtensorMap map_a1,
                  const __grid_constant__ CUtensorMap map_a2, const __grid_constant__ CUtensorMap map_a3,
                  const __grid_constant__ CUtensorMap map_w, const ConvKParams p) {
    using Cfg = ConvCfg<BN, STATS>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
    const uint32_t bar_base = smem_base + Cfg::kStages * Cfg::kStageBytes;
    // barrier slots (8 B each)
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (Cfg::kStages + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * Cfg::kStages + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * Cfg::kStages + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * Cfg::kStages + 4);
    volatile uint32_t* tmem_slot_ptr =
        reinterpret_cast<volatile uint32_t*>(smem_al + Cfg::kStages * Cfg::kStageBytes + 8 * (2 * Cfg::kStages + 4));
    float* s_scale = reinterpret_cast<float*>(smem_al + Cfg::kStages * Cfg::kStageBytes + Cfg::kBarBytes);
    float* s_bias = s_scale + kMaxCout;
    // STATS: per-CTA partial BatchNorm statistics [4 quarters][2][kStatMaxCout] fp32 + a 2 KB transpose scratch per epilogue
    // warp (vu_conv_common.cuh: epi_stats_chunk)
    float* s_stat = s_bias + kMaxCout;
    uint8_t* s_scratch = reinterpret_cast<uint8_t*>(s_stat) + kEpiStatBytes;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_a0);
        tma_prefetch_desc(&map_w);
        for (int s = 0; s < Cfg::kStages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 32 * kEpiWarps); }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc<Cfg::kTmemCols>(tmem_slot);
        tmem_relinquish();
    }
    for (int c = threadIdx.x; c < p.cout; c += kThreads) {
        s_scale[c] = p.scale ? p.scale[c] : 1.0f;
        s_bias[c] = p.bias ? p.bias[c] : 0.0f;
    }
    if constexpr (STATS)
        for (int c = threadIdx.x; c < kEpiStatBytes / 4; c += kThreads) s_stat[c] = 0.0f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    const int total_tiles = p.num_m_tiles * p.num_n_tiles;
    const int k_iters = p.num_taps * p.kblocks;

    if (warp == 0) {
        // ===================== TMA producer (one lane) =====================
        if (lane == 0) {
            const CUtensorMap* maps[4] = {&map_a0, &map_a1, &map_a2, &map_a3};
            // running stage / phase counters: no division or modulo in the per-k-iteration path of either pipeline end
            // (with 64-channel k-iterations and BN <= 128 the MMAs of a k-iteration take only ~256 clk)
            uint32_t s = 0, ph = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const int n_tile = tile % p.num_n_tiles;
                const int m_tile = tile / p.num_n_tiles;
                const int wb = m_tile % p.tiles_w;
                const int hb = (m_tile / p.tiles_w) % p.tiles_h;
                const int nb = m_tile / (p.tiles_w * p.tiles_h);
                const int w0 = wb * p.wt, h0 = hb * p.ht, n0 = nb * p.nt;
                for (int tap = 0; tap < p.num_taps; ++tap) {
                    const CUtensorMap* ma = maps[p.tap_map[tap]];
                    const int ww = w0 + p.tap_dw[tap], hh = h0 + p.tap_dh[tap];
                    const int widx = p.tap_widx[tap];
                    for (int kb = 0; kb < p.kblocks; ++kb) {
                        mbar_wait(empty_bar(s), ph ^ 1);
                        const uint32_t sa = smem_base + s * Cfg::kStageBytes;
                        mbar_arrive_expect_tx(full_bar(s), Cfg::kStageBytes);
                        tma_load_4d(sa, ma, full_bar(s), kb * kBlockK, ww, hh, n0);
                        tma_load_3d(sa + kABytes, &map_w, full_bar(s), kb * kBlockK, widx, n_tile * BN);
                        if (++s == Cfg::kStages) { s = 0; ph ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer: warp-uniform control flow, one elected lane issues =====
        // (elect.sync instead of `lane == 0`: ptxas then emits bare UTCHMMA, no per-MMA uniform-branch loop)
        const uint32_t idesc = umma_idesc_f16(kTileM, BN, p.dtype == 0 ? 1u : 0u);
        uint32_t s = 0, ph = 0, tcount = 0;
        const uint64_t desc0 = umma_desc_k_sw128(smem_base);          // stage 0; +stage bytes >> 4 per stage
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
            const int acc = tcount & 1;
            const uint32_t accph = (tcount >> 1) & 1;
            mbar_wait(tempty_bar(acc), accph ^ 1);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc * BN;
            int kb = 0;
            for (int ki = 0; ki < k_iters; ++ki) {
                mbar_wait(full_bar(s), ph);
                tc_fence_after();
                const uint64_t adesc = desc0 + (uint64_t)((s * Cfg::kStageBytes) >> 4);
                const uint64_t bdesc = adesc + (uint64_t)(kABytes >> 4);
                const bool last_kb = (kb == p.kblocks - 1);
                const int nk = last_kb ? p.last_ksteps : (kBlockK / 16);
                if (elect_one()) {
                    // +32 B per K=16 step inside the 128 B swizzle row (encoded >> 4)
                    if (nk == 4) {
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_f16_ss(tmem_d, adesc + 2u * k, bdesc + 2u * k, idesc, (ki | k) != 0);
                    } else {
                        for (int k = 0; k < nk; ++k)
                            umma_f16_ss(tmem_d, adesc + 2u * k, bdesc + 2u * k, idesc, (ki | k) != 0);
                    }
                    umma_commit(empty_bar(s));
                    if (ki == k_iters - 1) umma_commit(tfull_bar(acc));
                }
                __syncwarp();
                kb = last_kb ? 0 : kb + 1;
                if (++s == Cfg::kStages) { s = 0; ph ^= 1; }
            }
        }
    } else {
        // ===================== epilogue (8 warps: 2 per TMEM lane quarter) =====================
        const int q = warp & 3;                   // TMEM lane quarter this warp may access
        const int half = (warp - 2) >> 2;         // even / odd 32-column chunks
        uint8_t* scratch = s_scratch + (warp - 2) * kEpiScratchBytes;
        uint32_t tcount = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
            const int acc = tcount & 1;
            const uint32_t accph = (tcount >> 1) & 1;
            const int n_tile = tile % p.num_n_tiles;
            const int m_tile = tile / p.num_n_tiles;
            mbar_wait(tfull_bar(acc), accph);
            tc_fence_after();
            conv_epilogue_tile<BN, STATS>(p, tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN, n_tile, m_tile, q, half,
                                          lane, s_scale, s_bias, s_stat, scratch);
            tc_fence_before();
            mbar_arrive(tempty_bar(acc));
        }
    }

    tc_fence_before();
    __syncthreads();
    if constexpr (STATS)
        if (p.stat_sums != nullptr) epi_stats_flush(s_stat, p.stat_sums, p.cout, threadIdx.x, kThreads);
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<Cfg::kTmemCols>(tmem_base);
    }
}

// ------------------------------------------------------------------------------------------
// CTA-pair variant (cluster of 2, tcgen05 cta_group::2): one 256-pixel x BN tile per pair. Each CTA stages its own 128
// pixel rows of A and HALF of the filter rows of B; the leader's MMA reads both halves, so a k-iteration moves
// 16 KB + BN*64 B per SM through L2 instead of 16 KB + BN*128 B. The mid-resolution layers (K = 9*Cin deep, 128..512
// channels) sit on the L2->SM cap (~6300 B/clk chip-wide, profiles/r1a_conv_800to512_32sq.summary.txt), not on the tensor pipe.
//   full[s]   : leader's barrier, 2 arrivals (both producers) + the bytes of both CTAs' loads
//   empty[s]  : per CTA, released by the leader's MMA commit (multicast to both CTAs)
//   tfull[a]  : per CTA, same multicast commit; each CTA's epilogue drains its own TMEM lanes
//   tempty[a] : leader's barrier, 256 arrivals (the epilogue threads of both CTAs)
// ------------------------------------------------------------------------------------------
template <int BN, bool STATS> struct PairCfg {
    static constexpr int kBBytes = (BN / 2) * kBlockK * 2;          // this CTA's half of the filter rows
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kBarBytes = 1024;
    static constexpr int kTailBytes = kBarBytes + 2 * kMaxCout * 4 + (STATS ? kEpiStatBytes + kEpiWarps * kEpiScratchBytes : 0);
    static constexpr int kMaxStages = (kSmemLimit - 1024 - kTailBytes) / kStageBytes;
    static constexpr int kStages = kMaxStages > 8 ? 8 : kMaxStages;
    static constexpr int kTmemCols = 2 * BN;
    static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kTailBytes;
    static_assert(kStages >= 4, "pipeline depth");
};

template <int BN, bool STATS>
__global__ void __launch_bounds__(kThreads, 1)
conv_igemm_pair_kernel(const __grid_constant__ CUtensorMap map_a0, const __grid_constant__ CUtensorMap map_a1,
                       const __grid_constant__ CUtensorMap map_a2, const __grid_constant__ CUtensorMap map_a3,
                       const __grid_constant__ CUtensorMap map_w, const ConvKParams p) {
    using Cfg = PairCfg<BN, STATS>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
    const uint32_t bar_base = smem_base + Cfg::kStages * Cfg::kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (Cfg::kStages + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * Cfg::kStages + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * Cfg::kStages + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * Cfg::kStages + 4);
    volatile uint32_t* tmem_slot_ptr =
        reinterpret_cast<volatile uint32_t*>(smem_al + Cfg::kStages * Cfg::kStageBytes + 8 * (2 * Cfg::kStages + 4));
    float* s_scale = reinterpret_cast<float*>(smem_al + Cfg::kStages * Cfg::kStageBytes + Cfg::kBarBytes);
    float* s_bias = s_scale + kMaxCout;
    float* s_stat = s_bias + kMaxCout;
    uint8_t* s_scratch = reinterpret_cast<uint8_t*>(s_stat) + kEpiStatBytes;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();              // 0 = leader (issues the MMAs)
    const int pair_id = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_a0);
        tma_prefetch_desc(&map_w);
        for (int s = 0; s < Cfg::kStages; ++s) { mbar_init(full_bar(s), 2); mbar_init(empty_bar(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 2 * 32 * kEpiWarps); }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc_pair<Cfg::kTmemCols>(tmem_slot);
        tmem_relinquish_pair();
    }
    for (int c = threadIdx.x; c < p.cout; c += kThreads) {
        s_scale[c] = p.scale ? p.scale[c] : 1.0f;
        s_bias[c] = p.bias ? p.bias[c] : 0.0f;
    }
    if constexpr (STATS)
        for (int c = threadIdx.x; c < kEpiStatBytes / 4; c += kThreads) s_stat[c] = 0.0f;
    tc_fence_before();
    cluster_sync_all();                                   // barrier inits + TMEM allocation visible to the peer
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    const int m_pairs = (p.num_m_tiles + 1) >> 1;
    const int total_tiles = m_pairs * p.num_n_tiles;      // pair tiles
    const int k_iters = p.num_taps * p.kblocks;

    if (warp == 0) {
        // ===================== TMA producer (one lane per CTA) =====================
        if (lane == 0) {
            const CUtensorMap* maps[4] = {&map_a0, &map_a1, &map_a2, &map_a3};
            const uint32_t leader_full0 = mapa_cluster(full_bar(0), 0);
            uint32_t s = 0, ph = 0;
            for (int tile = pair_id; tile < total_tiles; tile += num_pairs) {
                const int n_tile = tile % p.num_n_tiles;
                const int m_tile = 2 * (tile / p.num_n_tiles) + (int)rank;   // may be one past the end: loads zero fill
                const int wb = m_tile % p.tiles_w;
                const int hb = (m_tile / p.tiles_w) % p.tiles_h;
                const int nb = m_tile / (p.tiles_w * p.tiles_h);
                const int w0 = wb * p.wt, h0 = hb * p.ht, n0 = nb * p.nt;
                for (int tap = 0; tap < p.num_taps; ++tap) {
                    const CUtensorMap* ma = maps[p.tap_map[tap]];
                    const int ww = w0 + p.tap_dw[tap], hh = h0 + p.tap_dh[tap];
                    const int widx = p.tap_widx[tap];
                    for (int kb = 0; kb < p.kblocks; ++kb) {
                        mbar_wait(empty_bar(s), ph ^ 1);
                        const uint32_t sa = smem_base + s * Cfg::kStageBytes;
                        if (rank == 0) mbar_arrive_expect_tx(full_bar(s), 2 * Cfg::kStageBytes);
                        else mbar_arrive_cluster(leader_full0 + 8u * s);
                        tma_load_4d_pair(sa, ma, full_bar(s), kb * kBlockK, ww, hh, n0);
                        tma_load_3d_pair(sa + kABytes, &map_w, full_bar(s), kb * kBlockK, widx,
                                         n_tile * BN + (int)rank * (BN / 2));
                        if (++s == Cfg::kStages) { s = 0; ph ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer: leader CTA only =====================
        if (rank == 0) {
            const uint32_t idesc = umma_idesc_f16(2 * kTileM, BN, p.dtype == 0 ? 1u : 0u);
            uint32_t s = 0, ph = 0, tcount = 0;
            const uint64_t desc0 = umma_desc_k_sw128(smem_base);
            for (int tile = pair_id; tile < total_tiles; tile += num_pairs, ++tcount) {
                const int acc = tcount & 1;
                const uint32_t accph = (tcount >> 1) & 1;
                mbar_wait(tempty_bar(acc), accph ^ 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                int kb = 0;
                for (int ki = 0; ki < k_iters; ++ki) {
                    mbar_wait(full_bar(s), ph);
                    tc_fence_after();
                    const uint64_t adesc = desc0 + (uint64_t)((s * Cfg::kStageBytes) >> 4);
                    const uint64_t bdesc = adesc + (uint64_t)(kABytes >> 4);
                    const bool last_kb = (kb == p.kblocks - 1);
                    const int nk = last_kb ? p.last_ksteps : (kBlockK / 16);
                    if (elect_one()) {
                        if (nk == 4) {
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                umma_f16_ss_pair(tmem_d, adesc + 2u * k, bdesc + 2u * k, idesc, (ki | k) != 0);
                        } else {
                            for (int k = 0; k < nk; ++k)
                                umma_f16_ss_pair(tmem_d, adesc + 2u * k, bdesc + 2u * k, idesc, (ki | k) != 0);
                        }
                        umma_commit_pair(empty_bar(s), 3);
                        if (ki == k_iters - 1) umma_commit_pair(tfull_bar(acc), 3);
                    }
                    __syncwarp();
                    kb = last_kb ? 0 : kb + 1;
                    if (++s == Cfg::kStages) { s = 0; ph ^= 1; }
                }
            }
        }
    } else {
        // ===================== epilogue: each CTA drains its own 128 TMEM lanes (8 warps) =====================
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        uint8_t* scratch = s_scratch + (warp - 2) * kEpiScratchBytes;
        const uint32_t leader_tempty0 = mapa_cluster(tempty_bar(0), 0);
        uint32_t tcount = 0;
        for (int tile = pair_id; tile < total_tiles; tile += num_pairs, ++tcount) {
            const int acc = tcount & 1;
            const uint32_t accph = (tcount >> 1) & 1;
            const int n_tile = tile % p.num_n_tiles;
            const int m_tile = 2 * (tile / p.num_n_tiles) + (int)rank;
            mbar_wait(tfull_bar(acc), accph);
            tc_fence_after();
            if (m_tile < p.num_m_tiles)
                conv_epilogue_tile<BN, STATS>(p, tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN, n_tile, m_tile, q,
                                              half, lane, s_scale, s_bias, s_stat, scratch);
            tc_fence_before();
            mbar_arrive_cluster(leader_tempty0 + 8u * acc);
        }
    }

    tc_fence_before();
    cluster_sync_all();                                   // no CTA leaves while the peer may still signal it / read its smem
    if constexpr (STATS)
        if (p.stat_sums != nullptr) epi_stats_flush(s_stat, p.stat_sums, p.cout, threadIdx.x, kThreads);
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
    }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static int pow2_ge(int v) { int p = 1; while (p < v) p <<= 1; return p; }

// choose (wt, ht, nt), product 128, maximising the fraction of useful rows
static void choose_tile(int n, int h, int w, int* wt_o, int* ht_o, int* nt_o) {
    double best = -1.0;
    for (int wt = 128; wt >= 1; wt >>= 1) {
        for (int ht = 128 / wt; ht >= 1; ht >>= 1) {
            const int nt = 128 / (wt * ht);
            if (nt > 128) continue;
            const double cover = (double)((w + wt - 1) / wt * wt) * ((h + ht - 1) / ht * ht) *
                                 ((n + nt - 1) / nt * nt);
            const double eff = (double)w * h * n / cover;
            if (eff > best + 1e-9) { best = eff; *wt_o = wt; *ht_o = ht; *nt_o = nt; }
        }
    }
}

template <int BN, bool STATS>
static int launch_conv(const CUtensorMap* am, const CUtensorMap& wm, const ConvKParams& kp, int max_ctas,
                       cudaStream_t stream) {
    using Cfg = ConvCfg<BN, STATS>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv_igemm_kernel<BN, STATS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::kSmemBytes);
        if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        attr_set = true;
    }
    int grid = kp.num_m_tiles * kp.num_n_tiles;
    const int sms = num_sms();
    int cap = (max_ctas > 0) ? max_ctas : sms;
    if (grid > cap) grid = cap;
    conv_igemm_kernel<BN, STATS><<<grid, kThreads, Cfg::kSmemBytes, stream>>>(am[0], am[1], am[2], am[3], wm, kp);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "conv_igemm launch: %s", cudaGetErrorString(e));
    return 0;
}

// CTA pairs: cluster of 2 along x. -> 0 launched, >0 error; `max_clusters` is how many pairs can be co-resident.
template <int BN, bool STATS>
static int launch_conv_pair(const CUtensorMap* am, const CUtensorMap& wm, const ConvKParams& kp, int max_ctas,
                            cudaStream_t stream) {
    using Cfg = PairCfg<BN, STATS>;
    static int max_clusters = -1;
    cudaLaunchConfig_t cfg{};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = Cfg::kSmemBytes; cfg.stream = stream;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (max_clusters < 0) {
        cudaError_t e = cudaFuncSetAttribute(conv_igemm_pair_kernel<BN, STATS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::kSmemBytes);
        if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        cfg.gridDim = dim3(2 * (num_sms() / 2));
        int n = 0;
        e = cudaOccupancyMaxActiveClusters(&n, conv_igemm_pair_kernel<BN, STATS>, &cfg);
        if (e != cudaSuccess || n < 1) return set_error(VU_ERR_CUDA, "cudaOccupancyMaxActiveClusters: %s", cudaGetErrorString(e));
        max_clusters = n;
    }
    int pairs = ((kp.num_m_tiles + 1) / 2) * kp.num_n_tiles;
    int cap = max_clusters;
    if (max_ctas > 0 && max_ctas / 2 < cap) cap = max_ctas / 2 > 0 ? max_ctas / 2 : 1;
    if (pairs > cap) pairs = cap;
    cfg.gridDim = dim3(2 * pairs);
    cudaError_t e = cudaLaunchKernelEx(&cfg, conv_igemm_pair_kernel<BN, STATS>, am[0], am[1], am[2], am[3], wm, kp);
    if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "conv_igemm_pair launch: %s", cudaGetErrorString(e));
    return 0;
}

}  // namespace vu

using namespace vu;

extern "C" int vu_conv2d_fwd(const vu_conv_args* a, vu_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!a || !a->x || !a->w || (!a->y && !a->fuse_out)) return set_error(VU_ERR_INVALID, "conv: null pointer");
    const int KH = a->KH, KW = a->KW, S = a->stride, P = a->pad;
    const bool ok_geom = (KH == KW) && ((KH == 3 && P == 1) || (KH == 1 && P == 0)) && (S == 1 || S == 2);
    if (!ok_geom) return set_error(VU_ERR_UNSUPPORTED, "conv: only 3x3 p1 / 1x1 p0, stride 1|2 (got k%d s%d p%d)", KH, S, P);
    if (a->Cin < 8 || a->Cin % 8 || a->x_ctot % 8 || a->x_coff % 8 || a->y_ctot % 8 || a->y_coff % 8)
        return set_error(VU_ERR_ALIGN, "conv: channel counts/offsets must be multiples of 8 (Cin=%d x_ctot=%d x_coff=%d y_ctot=%d y_coff=%d)",
                         a->Cin, a->x_ctot, a->x_coff, a->y_ctot, a->y_coff);
    if (a->residual && (a->res_ctot % 8 || a->res_coff % 8)) return set_error(VU_ERR_ALIGN, "conv: residual slice misaligned");
    if (a->Cout % 32 || a->Cout > kMaxCout) return set_error(VU_ERR_UNSUPPORTED, "conv: Cout must be a multiple of 32 and <= %d (got %d)", kMaxCout, a->Cout);
    if ((reinterpret_cast<uintptr_t>(a->x) | reinterpret_cast<uintptr_t>(a->w) | reinterpret_cast<uintptr_t>(a->y) |
         reinterpret_cast<uintptr_t>(a->residual)) & 15)
        return set_error(VU_ERR_ALIGN, "conv: pointers must be 16-byte aligned");
    if (a->N < 1 || a->H < 1 || a->W < 1) return set_error(VU_ERR_INVALID, "conv: empty input");
    const int up = a->out_up > 0 ? a->out_up : 1;
    {   // small-Cout 3x3 on wide maps: the smem tap-reuse kernel (vu_conv_halo.cu)
        const int rc = conv3x3_halo_launch(a, stream);
        if (rc != 0) return rc < 0 ? rc : 0;
    }
    {   // 3x3 on narrow maps (W <= 64): tap reuse over a flat padded pixel index (vu_conv_flat.cu)
        const int rc = conv3x3_flat_launch(a, stream);
        if (rc != 0) return rc < 0 ? rc : 0;
    }
    if (a->fuse_out || !a->y)
        return set_error(VU_ERR_UNSUPPORTED, "conv: fused output conv needs 3x3 s1 p1, Cout 64/128, W >= 96");

    const int Ho = (a->H + 2 * P - KH) / S + 1, Wo = (a->W + 2 * P - KW) / S + 1;
    ConvKParams kp{};
    kp.n_img = a->N; kp.ho = Ho; kp.wo = Wo;
    choose_tile(a->N, Ho, Wo, &kp.wt, &kp.ht, &kp.nt);
    kp.tiles_w = (Wo + kp.wt - 1) / kp.wt; kp.tiles_h = (Ho + kp.ht - 1) / kp.ht; kp.tiles_n = (a->N + kp.nt - 1) / kp.nt;
    kp.num_m_tiles = kp.tiles_w * kp.tiles_h * kp.tiles_n;
    kp.cout = a->Cout;
    kp.num_taps = KH * KW;
    kp.kblocks = (a->Cin + kBlockK - 1) / kBlockK;
    kp.last_ksteps = ((a->Cin - (kp.kblocks - 1) * kBlockK) + 15) / 16;
    kp.scale = a->scale; kp.bias = a->bias;
    kp.residual = static_cast<const uint16_t*>(a->residual); kp.res_ctot = a->res_ctot; kp.res_coff = a->res_coff;
    kp.y = static_cast<uint16_t*>(a->y); kp.y_ctot = a->y_ctot; kp.y_coff = a->y_coff;
    kp.out_up = up; kp.out_dh = a->out_dh; kp.out_dw = a->out_dw;
    kp.out_h = Ho * up; kp.out_w = Wo * up;
    kp.relu = a->relu; kp.dtype = a->dtype;
    kp.pos_bias = a->pos_bias;
    kp.stat_sums = a->stat_sums;
    if (a->stat_sums && up != 1) return set_error(VU_ERR_UNSUPPORTED, "conv: stat_sums with the pixel-shuffle store");
    if (a->pos_bias && !(KH == 3 && KW == 3 && S == 1 && P == 1 && up == 1 && a->H >= 2 && a->W >= 2))
        return set_error(VU_ERR_UNSUPPORTED, "conv: pos_bias needs a 3x3 s1 p1 conv on a map of at least 2x2");

    // BN tile of 256/128/64/32 columns; the last N tile may be partial (filter rows beyond Cout are TMA zero fill, the
    // epilogue skips them). A k-iteration is L2->SM bound, so pick the width that moves the fewest operand rows:
    // n_tiles * (128 pixel rows + BN filter rows); ties go to the narrower tile (fewer wasted MMA columns).
    int bn = 32;
    long best = -1;
    for (int cand = 32; cand <= 256; cand <<= 1) {
        const long cost = (long)((a->Cout + cand - 1) / cand) * (128 + cand);
        if (best < 0 || cost < best) { best = cost; bn = cand; }
    }
    const int sms = num_sms();
    // only drop from 256 to 128 columns when the wide tiles would leave more than ~40 % of the SMs idle: a single
    // 86 %-full wave of 256-column tiles moves fewer bytes through L2 than two rounds of 128-column tiles
    while (bn > 128 && (long)kp.num_m_tiles * ((a->Cout + bn - 1) / bn) * 5 < (long)sms * 3) bn >>= 1;
    kp.num_n_tiles = (a->Cout + bn - 1) / bn;

    // activation tensor maps
    CUtensorMap am[4];
    const uint16_t* xb = static_cast<const uint16_t*>(a->x) + a->x_coff;
    const uint32_t box[4] = {(uint32_t)kBlockK, (uint32_t)kp.wt, (uint32_t)kp.ht, (uint32_t)kp.nt};
    const uint64_t pix = (uint64_t)a->x_ctot * 2;
    int nmaps = 0;
    if (S == 1) {
        const uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->N};
        const uint64_t str[3] = {pix, pix * a->W, pix * a->W * a->H};
        int rc = encode_map(&am[0], a->dtype, 4, xb, dims, str, box);
        if (rc) return rc;
        nmaps = 1;
        int t = 0;
        for (int r = 0; r < KH; ++r)
            for (int s = 0; s < KW; ++s) {
                if (a->tap_mask != 0 && !((a->tap_mask >> (r * KW + s)) & 1)) continue;   // all-zero filter tap
                kp.tap_dh[t] = (int8_t)(r - P); kp.tap_dw[t] = (int8_t)(s - P);
                kp.tap_map[t] = 0; kp.tap_widx[t] = (int8_t)(r * KW + s);
                ++t;
            }
        if (t == 0) return set_error(VU_ERR_INVALID, "conv: tap_mask selects no tap");
        kp.num_taps = t;
    } else {
        if (a->tap_mask != 0) return set_error(VU_ERR_UNSUPPORTED, "conv: tap_mask needs stride 1");
        // parity planes: plane (pr,pc) holds rows pr, pr+2, ... and cols pc, pc+2, ...
        for (int pr = 0; pr < 2; ++pr)
            for (int pc = 0; pc < 2; ++pc) {
                const int hp = (a->H - pr + 1) / 2, wp = (a->W - pc + 1) / 2;
                const int id = pr * 2 + pc;
                if (hp <= 0 || wp <= 0) { am[id] = am[0]; continue; }
                const uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)wp, (uint64_t)hp, (uint64_t)a->N};
                const uint64_t str[3] = {pix * 2, pix * a->W * 2, pix * a->W * a->H};
                int rc = encode_map(&am[id], a->dtype, 4, xb + ((size_t)pr * a->W + pc) * a->x_ctot, dims, str, box);
                if (rc) return rc;
            }
        nmaps = 4;
        int t = 0;
        for (int r = 0; r < KH; ++r)
            for (int s = 0; s < KW; ++s, ++t) {
                const int ir = r - P, ic = s - P;             // input offset relative to 2*ho, 2*wo
                const int pr = ((ir % 2) + 2) % 2, pc = ((ic % 2) + 2) % 2;
                kp.tap_dh[t] = (int8_t)((ir - pr) / 2); kp.tap_dw[t] = (int8_t)((ic - pc) / 2);
                kp.tap_map[t] = (int8_t)(pr * 2 + pc); kp.tap_widx[t] = (int8_t)t;
            }
    }
    for (int i = nmaps; i < 4; ++i) am[i] = am[0];

    // CTA pairs (256-pixel tiles, half of the filter rows per SM) for the wide tiles: fewer operand bytes per MMA through L2
    static const int pair_mode = getenv("VU_CONV_PAIR") ? atoi(getenv("VU_CONV_PAIR")) : 1;
    const bool pair = pair_mode != 0 && bn >= 128 && kp.num_m_tiles >= 2;

    // weight map: [Cout][taps][Cin]
    CUtensorMap wm;
    {
        const uint64_t dims[3] = {(uint64_t)a->Cin, (uint64_t)(KH * KW), (uint64_t)a->Cout};
        const uint64_t str[2] = {(uint64_t)a->Cin * 2, (uint64_t)a->Cin * 2 * KH * KW};
        const uint32_t wbox[3] = {(uint32_t)kBlockK, 1u, (uint32_t)(pair ? bn / 2 : bn)};
        int rc = encode_map(&wm, a->dtype, 3, a->w, dims, str, wbox);
        if (rc) return rc;
    }
    const bool stats = a->stat_sums != nullptr;
    if (stats && a->Cout > kStatMaxCout)
        return set_error(VU_ERR_UNSUPPORTED, "conv: stat_sums needs Cout <= %d (got %d): use vu_bn_stats", kStatMaxCout, a->Cout);
    if (pair) {
        if (stats) return bn == 256 ? launch_conv_pair<256, true>(am, wm, kp, a->max_ctas, stream)
                                    : launch_conv_pair<128, true>(am, wm, kp, a->max_ctas, stream);
        return bn == 256 ? launch_conv_pair<256, false>(am, wm, kp, a->max_ctas, stream)
                         : launch_conv_pair<128, false>(am, wm, kp, a->max_ctas, stream);
    }
    if (stats) {
        switch (bn) {
            case 256: return launch_conv<256, true>(am, wm, kp, a->max_ctas, stream);
            case 128: return launch_conv<128, true>(am, wm, kp, a->max_ctas, stream);
            case 64: return launch_conv<64, true>(am, wm, kp, a->max_ctas, stream);
            default: return launch_conv<32, true>(am, wm, kp, a->max_ctas, stream);
        }
    }
    switch (bn) {
        case 256: return launch_conv<256, false>(am, wm, kp, a->max_ctas, stream);
        case 128: return launch_conv<128, false>(am, wm, kp, a->max_ctas, stream);
        case 64: return launch_conv<64, false>(am, wm, kp, a->max_ctas, stream);
        default: return launch_conv<32, false>(am, wm, kp, a->max_ctas, stream);
    }
}
