// 3x3 s1 p1 convolution for small Cout (64 / 128) on wide maps — tcgen05 implicit GEMM with the
// nine taps served from ONE shared-memory copy of the input tile.
//
// Why a second kernel: conv_igemm_kernel fetches a fresh 128-pixel x 64-channel A tile from L2 for
// every tap, i.e. every input pixel crosses the L2->SM crossbar 9 times. With Cout = 64 an A element
// feeds only 64 MACs, so that kernel is bound by L2->SM bandwidth (profiles/r1a_conv_224to64_256sq:
// 8.5 TB/s, tensor pipe 15 %). Here a CTA owns an output tile of HT rows x 128 pixels; per 64-channel
// K block it loads the (HT+2) x 130-pixel halo'd input tile ONCE (one 4-D TMA box, zero padding = TMA
// out-of-bounds fill) and issues the MMAs of all nine taps from it: the A operand of tap (r,s) for
// output row j is the same smem tile with the descriptor start address advanced by
// ((j+r)*130 + s) * 128 bytes. Rows are 128 B (64 channels, SWIZZLE_128B); the swizzle is a function of
// the absolute smem address, so a start address that is 128-byte- but not 1024-byte-aligned still
// reads the rows TMA wrote. L2->SM traffic per MMA drops 3x (HT=2) to 5x (HT=4).
//
//   GEMM view per tile:  HT accumulators [128 pixels x BN], BN = Cout, K = 9 taps x Cin.
//   Warps:  0 = A (input tile) TMA producer, 1 = TMEM allocator + MMA issuer, 2 = B (filter tap) TMA
//           producer, 3..10 = epilogue (TMEM -> scale/bias/residual/ReLU -> 16-bit NHWC channel slice): two warps per
//           TMEM lane quarter, taking the even / odd output rows of a tile (one warp per scheduler could not drain a
//           4-row x 64-channel tile in the ~5k clk its MMAs take: ncu r2a, tensor pipe 16 % on the 64 -> 64 @128^2 layers).
//   Rings:  A: 2 stages of (HT+2)*130 rows; B: one [BN x 64ch] filter tap per stage; TMEM: 2 sets of HT
//           accumulators so the epilogue of tile i overlaps the MMAs of tile i+1.
#include "vu_conv_common.cuh"
#include <cstdlib>

namespace vu {

constexpr int kHaloWT = 128;               // output pixels per tile row = UMMA M
constexpr int kHaloPitch = kHaloWT + 2;    // input pixels per tile row (1-pixel halo each side)
constexpr int kHaloEpiWarps = 8;
constexpr int kHaloThreads = 96 + 32 * kHaloEpiWarps;

struct HaloParams {
    int n_img, h, w;
    int tiles_w, tiles_h, num_tiles;
    int kblocks, last_ksteps;
    const float* scale; const float* bias;
    const uint16_t* residual; int res_ctot, res_coff;
    uint16_t* y; int y_ctot, y_coff;
    int relu, dtype;
    const float* fuse_w; float fuse_bias; float* fuse_out;   // optional fused 1x1 Cout->1 output conv (fp32 [N,H,W])
    const float* pos_bias;                                    // optional [N][9][BN] pre-scale bias by border class
    double* stat_sums;                                        // optional [2*BN] fp64 sum / sum of squares of the stored outputs
};

// PAIR: cluster of two CTAs, tcgen05 cta_group::2 - each CTA stages its own pixel tile and HALF of the filter rows of a tap;
// the leader's M = 256 MMA reads both (see conv_igemm_pair_kernel). Per MMA an SM then reads 4 KB + BN*16 B of smem operands
// instead of 4 KB + BN*32 B: the Cout = 64 layers are bound by that read port.
// STATS: the training flavour that accumulates BatchNorm statistics of the stored outputs (per-CTA partials
// [8 epilogue warps][2][BN] fp32 - here the two warps of a lane quarter take different ROWS of the same channels, so every
// warp needs its own slice - + a 2 KB transpose scratch per epilogue warp, vu_conv_common.cuh); it runs with HT = 2.
template <int BN, int HT, bool PAIR = false, bool STATS = false> struct HaloCfg {
    static constexpr int kARows = (HT + 2) * kHaloPitch;
    static constexpr int kABytes = kARows * 128;                       // what one TMA box delivers
    static constexpr int kAStride = (kABytes + 1023) / 1024 * 1024;   // stage pitch (1024-aligned: swizzle atom)
    static constexpr int kAStages = 2;
    static constexpr int kBBytes = (PAIR ? BN / 2 : BN) * 128;
    static constexpr int kBStages = (BN == 64) ? (HT == 2 ? 8 : 3) : ((STATS && !PAIR) ? 3 : 5);
    static constexpr int kTmemCols = 2 * HT * BN;
    static constexpr int kBarBytes = 256;
    static constexpr int kStatBytes = STATS ? (kHaloEpiWarps * 2 * BN * 4 + kHaloEpiWarps * kEpiScratchBytes) : 0;
    static constexpr int kSmemBytes = 1024 + kAStages * kAStride + kBStages * kBBytes + kBarBytes + 3 * BN * 4 + kStatBytes;
    static_assert(kTmemCols <= 512, "TMEM budget");
    static_assert(kSmemBytes <= 232448, "smem budget");
    static_assert(!STATS || HT == 2, "the statistics flavour runs with two rows per tile");
};

template <int BN, int HT, bool PAIR, bool STATS>
__global__ void __launch_bounds__(kHaloThreads, 1)
conv3x3_halo_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                    const HaloParams p) {
    using Cfg = HaloCfg<BN, HT, PAIR, STATS>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
    const uint32_t a_base = smem_base;
    const uint32_t b_base = smem_base + Cfg::kAStages * Cfg::kAStride;
    constexpr int kBarOff = Cfg::kAStages * Cfg::kAStride + Cfg::kBStages * Cfg::kBBytes;
    const uint32_t bar_base = smem_base + kBarOff;
    auto a_full = [&](int s) { return bar_base + 8u * s; };
    auto a_empty = [&](int s) { return bar_base + 8u * (Cfg::kAStages + s); };
    auto b_full = [&](int s) { return bar_base + 8u * (2 * Cfg::kAStages + s); };
    auto b_empty = [&](int s) { return bar_base + 8u * (2 * Cfg::kAStages + Cfg::kBStages + s); };
    auto t_full = [&](int a) { return bar_base + 8u * (2 * Cfg::kAStages + 2 * Cfg::kBStages + a); };
    auto t_empty = [&](int a) { return bar_base + 8u * (2 * Cfg::kAStages + 2 * Cfg::kBStages + 2 + a); };
    constexpr int kSlotIdx = 2 * Cfg::kAStages + 2 * Cfg::kBStages + 4;
    static_assert(8 * (kSlotIdx + 1) <= Cfg::kBarBytes, "barrier area");
    const uint32_t tmem_slot = bar_base + 8u * kSlotIdx;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_al + kBarOff + 8 * kSlotIdx);
    float* s_scale = reinterpret_cast<float*>(smem_al + kBarOff + Cfg::kBarBytes);
    float* s_bias = s_scale + BN;
    float* s_fw = s_bias + BN;
    float* s_stat = s_fw + BN;                              // STATS: [8 epilogue warps][2][BN] fp32 partial BatchNorm statistics
    uint8_t* s_scratch = reinterpret_cast<uint8_t*>(s_stat + kHaloEpiWarps * 2 * BN);   // + 2 KB transpose scratch per warp

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    // PAIR: work items are pairs of tiles (2t, 2t+1); full barriers live in the leader (rank 0) and take both producers,
    // empty / accumulator-full barriers are per CTA (multicast commit), accumulator-empty lives in the leader
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
    const int item0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int item_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    const int num_items = PAIR ? (p.num_tiles + 1) / 2 : p.num_tiles;
    auto tile_of = [&](int item) { return PAIR ? 2 * item + (int)rank : item; };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_w);
        for (int s = 0; s < Cfg::kAStages; ++s) { mbar_init(a_full(s), PAIR ? 2 : 1); mbar_init(a_empty(s), 1); }
        for (int s = 0; s < Cfg::kBStages; ++s) { mbar_init(b_full(s), PAIR ? 2 : 1); mbar_init(b_empty(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(t_full(a), 1); mbar_init(t_empty(a), (PAIR ? 2 : 1) * 32 * kHaloEpiWarps); }
        fence_barrier_init();
    }
    if (warp == 1) {
        if constexpr (PAIR) { tmem_alloc_pair<Cfg::kTmemCols>(tmem_slot); tmem_relinquish_pair(); }
        else { tmem_alloc<Cfg::kTmemCols>(tmem_slot); tmem_relinquish(); }
    }
    for (int c = threadIdx.x; c < BN; c += kHaloThreads) {
        s_scale[c] = p.scale ? p.scale[c] : 1.0f;
        s_bias[c] = p.bias ? p.bias[c] : 0.0f;
        s_fw[c] = p.fuse_w ? p.fuse_w[c] : 0.0f;
    }
    if constexpr (STATS)
        for (int c = threadIdx.x; c < kHaloEpiWarps * 2 * BN; c += kHaloThreads) s_stat[c] = 0.0f;
    tc_fence_before();
    if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        // ===================== A producer: one halo'd input tile per (tile, K block) =====================
        if (lane == 0) {
            uint32_t it = 0;
            const uint32_t leader_a_full0 = PAIR ? mapa_cluster(a_full(0), 0) : 0u;
            for (int item = item0; item < num_items; item += item_step) {
                const int tile = tile_of(item);               // PAIR: may be one past the end (loads are zero fill)
                const int wb = tile % p.tiles_w;
                const int hb = (tile / p.tiles_w) % p.tiles_h;
                const int nb = tile / (p.tiles_w * p.tiles_h);
                for (int kb = 0; kb < p.kblocks; ++kb, ++it) {
                    const int s = it % Cfg::kAStages;
                    const uint32_t ph = (it / Cfg::kAStages) & 1;
                    mbar_wait(a_empty(s), ph ^ 1);
                    if (!PAIR) mbar_arrive_expect_tx(a_full(s), Cfg::kABytes);
                    else if (rank == 0) mbar_arrive_expect_tx(a_full(s), 2 * Cfg::kABytes);
                    else mbar_arrive_cluster(leader_a_full0 + 8u * s);
                    // one TMA op per tile row: several ops in flight fill the tile ~4x faster than one big box
#pragma unroll
                    for (int i = 0; i < HT + 2; ++i) {
                        if constexpr (PAIR)
                            tma_load_4d_pair(a_base + s * Cfg::kAStride + i * (kHaloPitch * 128), &map_x, a_full(s), kb * 64,
                                             wb * kHaloWT - 1, hb * HT - 1 + i, nb);
                        else
                            tma_load_4d(a_base + s * Cfg::kAStride + i * (kHaloPitch * 128), &map_x, a_full(s), kb * 64,
                                        wb * kHaloWT - 1, hb * HT - 1 + i, nb);
                    }
                }
            }
        }
    } else if (warp == 2) {
        // ===================== B producer: one [BN x 64ch] filter tap per stage =====================
        if (lane == 0) {
            uint32_t it = 0;
            const uint32_t leader_b_full0 = PAIR ? mapa_cluster(b_full(0), 0) : 0u;
            for (int item = item0; item < num_items; item += item_step) {
                for (int kb = 0; kb < p.kblocks; ++kb) {
                    for (int tap = 0; tap < 9; ++tap, ++it) {
                        const int s = it % Cfg::kBStages;
                        const uint32_t ph = (it / Cfg::kBStages) & 1;
                        mbar_wait(b_empty(s), ph ^ 1);
                        if constexpr (PAIR) {
                            if (rank == 0) mbar_arrive_expect_tx(b_full(s), 2 * Cfg::kBBytes);
                            else mbar_arrive_cluster(leader_b_full0 + 8u * s);
                            tma_load_3d_pair(b_base + s * Cfg::kBBytes, &map_w, b_full(s), kb * 64, tap, (int)rank * (BN / 2));
                        } else {
                            mbar_arrive_expect_tx(b_full(s), Cfg::kBBytes);
                            tma_load_3d(b_base + s * Cfg::kBBytes, &map_w, b_full(s), kb * 64, tap, 0);
                        }
                    }
                }
            }
        }
    } else if (warp == 1 && (!PAIR || rank == 0)) {
        // ===================== MMA issuer: the whole warp walks the pipeline, one elected lane issues =====
        // (issuing under elect.sync instead of `lane == 0` lets ptxas emit bare UTCHMMA without a per-MMA
        //  uniform-branch loop; the tap loop is unrolled so descriptor offsets are immediates)
        const uint32_t idesc = umma_idesc_f16(PAIR ? 256 : 128, BN, p.dtype == 0 ? 1u : 0u);
        auto mma = [&](uint32_t d, uint64_t a, uint64_t b, uint32_t acc_flag) {
            if constexpr (PAIR) umma_f16_ss_pair(d, a, b, idesc, acc_flag); else umma_f16_ss(d, a, b, idesc, acc_flag);
        };
        auto commit = [&](uint32_t bar) {
            if constexpr (PAIR) umma_commit_pair(bar, 3); else umma_commit(bar);
        };
        uint32_t ita = 0, itb = 0, tcount = 0;
        for (int item = item0; item < num_items; item += item_step, ++tcount) {
            const int acc = tcount & 1;
            const uint32_t accph = (tcount >> 1) & 1;
            mbar_wait(t_empty(acc), accph ^ 1);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc * (HT * BN);
            for (int kb = 0; kb < p.kblocks; ++kb, ++ita) {
                const int sa = ita % Cfg::kAStages;
                mbar_wait(a_full(sa), (ita / Cfg::kAStages) & 1);
                tc_fence_after();
                const uint64_t a0 = umma_desc_k_sw128(a_base + sa * Cfg::kAStride);
                const int nk = (kb == p.kblocks - 1) ? p.last_ksteps : 4;
#pragma unroll
                for (int tap = 0; tap < 9; ++tap, ++itb) {
                    const int sb = itb % Cfg::kBStages;
                    mbar_wait(b_full(sb), (itb / Cfg::kBStages) & 1);
                    tc_fence_after();
                    const uint64_t bdesc = umma_desc_k_sw128(b_base + sb * Cfg::kBBytes);
                    const int r = tap / 3, s = tap - 3 * r;
                    if (elect_one()) {
                        // k outer, rows inner: consecutive MMAs go to different accumulators. Tap (r,s) of output
                        // row j = the same smem tile, start advanced by ((j+r)*130+s) pixel rows of 128 B (>>4 = *8).
                        if (nk == 4) {
#pragma unroll
                            for (int k = 0; k < 4; ++k)
#pragma unroll
                                for (int j = 0; j < HT; ++j)
                                    mma(tmem_d + j * BN, a0 + (uint32_t)(((j + r) * kHaloPitch + s) * 8 + 2 * k),
                                        bdesc + 2u * k, (kb | tap | k) != 0);
                        } else {
                            for (int k = 0; k < nk; ++k)
#pragma unroll
                                for (int j = 0; j < HT; ++j)
                                    mma(tmem_d + j * BN, a0 + (uint32_t)(((j + r) * kHaloPitch + s) * 8 + 2 * k),
                                        bdesc + 2u * k, (kb | tap | k) != 0);
                        }
                        commit(b_empty(sb));
                        if (tap == 8) commit(a_empty(sa));
                    }
                    __syncwarp();
                }
            }
            if (elect_one()) commit(t_full(acc));
            __syncwarp();
        }
    } else if (warp >= 3) {
        // ===================== epilogue: warps 3..10, two per TMEM lane quarter (even / odd rows of the tile) ==========
        const int q = warp & 3;
        const int half = (warp - 3) >> 2;
        const int m = q * 32 + lane;              // pixel within the tile row
        const bool has_res = p.residual != nullptr, relu = p.relu != 0, store_y = p.y != nullptr;
        const bool fuse = p.fuse_out != nullptr;  // 1x1 Cout->1 output conv on the fp32 activations (final_conv)
        const bool affine = p.scale != nullptr || p.bias != nullptr;
        const int dtype = p.dtype;
        uint8_t* scratch = s_scratch + (warp - 3) * kEpiScratchBytes;
        uint32_t tcount = 0;
        const uint32_t leader_t_empty0 = PAIR ? mapa_cluster(t_empty(0), 0) : 0u;
        for (int item = item0; item < num_items; item += item_step, ++tcount) {
            const int tile = tile_of(item);
            const int acc = tcount & 1;
            const uint32_t accph = (tcount >> 1) & 1;
            const int wb = tile % p.tiles_w;
            const int hb = (tile / p.tiles_w) % p.tiles_h;
            const int nb = tile / (p.tiles_w * p.tiles_h);
            const int wo = wb * kHaloWT + m;
            mbar_wait(t_full(acc), accph);
            tc_fence_after();
#pragma unroll 1
            for (int j = half; j < (tile < p.num_tiles ? HT : 0); j += 2) {   // PAIR: the peer's tile past the end is empty
                const int ho = hb * HT + j;
                const bool valid = (wo < p.w) && (ho < p.h);
                const size_t pix = (static_cast<size_t>(nb) * p.h + ho) * p.w + wo;
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * (HT * BN) + j * BN;
                float dot = 0.0f;
#pragma unroll 1
                for (int c = 0; c < BN / 32; ++c) {
                    uint32_t r[32];
                    tmem_ld_32x32b_x32(taddr + c * 32, r);
                    const int ch0 = c * 32;
                    uint4 rv[4];
                    if (has_res) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) rv[i] = make_uint4(0, 0, 0, 0);
                        if (valid) {
                            const uint4* rp = reinterpret_cast<const uint4*>(p.residual + pix * p.res_ctot + p.res_coff + ch0);
#pragma unroll
                            for (int i = 0; i < 4; ++i) rv[i] = __ldg(rp + i);
                        }
                    }
                    tmem_ld_wait();
                    if (p.pos_bias != nullptr) {
                        const int hc = min(ho, p.h - 1), wc = min(wo, p.w - 1);
                        const int cls = (hc == 0 ? 0 : (hc == p.h - 1 ? 2 : 1)) * 3 + (wc == 0 ? 0 : (wc == p.w - 1 ? 2 : 1));
                        const float4* pp = reinterpret_cast<const float4*>(p.pos_bias + (static_cast<size_t>(nb) * 9 + cls) * BN + ch0);
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float4 b = __ldg(pp + i);
                            r[4 * i] = __float_as_uint(__uint_as_float(r[4 * i]) + b.x);
                            r[4 * i + 1] = __float_as_uint(__uint_as_float(r[4 * i + 1]) + b.y);
                            r[4 * i + 2] = __float_as_uint(__uint_as_float(r[4 * i + 2]) + b.z);
                            r[4 * i + 3] = __float_as_uint(__uint_as_float(r[4 * i + 3]) + b.w);
                        }
                    }
                    float v[32];
                    if (affine) {
                        const float4* sc4 = reinterpret_cast<const float4*>(s_scale + ch0);
                        const float4* sb4 = reinterpret_cast<const float4*>(s_bias + ch0);
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float4 a = sc4[i], b = sb4[i];
                            v[4 * i] = fmaf(__uint_as_float(r[4 * i]), a.x, b.x);
                            v[4 * i + 1] = fmaf(__uint_as_float(r[4 * i + 1]), a.y, b.y);
                            v[4 * i + 2] = fmaf(__uint_as_float(r[4 * i + 2]), a.z, b.z);
                            v[4 * i + 3] = fmaf(__uint_as_float(r[4 * i + 3]), a.w, b.w);
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
                    }
                    if (has_res) {
                        const uint32_t* rw = reinterpret_cast<const uint32_t*>(rv);
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const float2 rr = unpack2(rw[i], dtype);
                            v[2 * i] += rr.x; v[2 * i + 1] += rr.y;
                        }
                    }
                    if (relu) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.0f);
                    }
                    if (fuse) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) dot = fmaf(v[i], s_fw[ch0 + i], dot);
                    }
                    if (store_y) {
                        uint32_t outw[16];
                        if (dtype == 0) {
#pragma unroll
                            for (int i = 0; i < 16; ++i) outw[i] = pack2(v[2 * i], v[2 * i + 1], 0);
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; ++i) outw[i] = pack2(v[2 * i], v[2 * i + 1], 1);
                        }
                        if (valid) {
                            uint4* op = reinterpret_cast<uint4*>(p.y + pix * p.y_ctot + p.y_coff + ch0);
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                op[i] = make_uint4(outw[4 * i], outw[4 * i + 1], outw[4 * i + 2], outw[4 * i + 3]);
                        }
                        if constexpr (STATS) {
                            if (p.stat_sums != nullptr)      // statistics of the values as stored (rounded to 16 bit)
                                epi_stats_chunk(outw, valid, lane, dtype, scratch, s_stat + ((warp - 3) * 2) * BN + ch0,
                                                s_stat + ((warp - 3) * 2 + 1) * BN + ch0);
                        }
                    }
                }
                if (fuse && valid) p.fuse_out[pix] = dot + p.fuse_bias;
            }
            tc_fence_before();
            if constexpr (PAIR) mbar_arrive_cluster(leader_t_empty0 + 8u * acc); else mbar_arrive(t_empty(acc));
        }
    }

    tc_fence_before();
    if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
    if constexpr (STATS)
        if (p.stat_sums != nullptr) epi_stats_flush(s_stat, p.stat_sums, BN, threadIdx.x, kHaloThreads, BN, kHaloEpiWarps);
    if (warp == 1) {
        tc_fence_after();
        if constexpr (PAIR) tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base); else tmem_dealloc<Cfg::kTmemCols>(tmem_base);
    }
}

template <int BN, int HT, bool STATS>
static int launch_halo(const CUtensorMap& xm, const CUtensorMap& wm, const HaloParams& hp, int max_ctas,
                       cudaStream_t stream) {
    using Cfg = HaloCfg<BN, HT, false, STATS>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv3x3_halo_kernel<BN, HT, false, STATS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::kSmemBytes);
        if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "cudaFuncSetAttribute(halo): %s", cudaGetErrorString(e));
        attr_set = true;
    }
    int grid = hp.num_tiles;
    const int cap = (max_ctas > 0) ? max_ctas : num_sms();
    if (grid > cap) grid = cap;
    conv3x3_halo_kernel<BN, HT, false, STATS><<<grid, kHaloThreads, Cfg::kSmemBytes, stream>>>(xm, wm, hp);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "conv3x3_halo launch: %s", cudaGetErrorString(e));
    return 1;
}

// CTA pairs (cluster of 2 along x); `wm` must have a box of BN/2 filter rows.
template <int BN, int HT, bool STATS>
static int launch_halo_pair(const CUtensorMap& xm, const CUtensorMap& wm, const HaloParams& hp, int max_ctas,
                            cudaStream_t stream) {
    using Cfg = HaloCfg<BN, HT, true, STATS>;
    static int max_clusters = -1;
    cudaLaunchConfig_t cfg{};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.blockDim = dim3(kHaloThreads); cfg.dynamicSmemBytes = Cfg::kSmemBytes; cfg.stream = stream;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (max_clusters < 0) {
        cudaError_t e = cudaFuncSetAttribute(conv3x3_halo_kernel<BN, HT, true, STATS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::kSmemBytes);
        if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "cudaFuncSetAttribute(halo pair): %s", cudaGetErrorString(e));
        cfg.gridDim = dim3(2 * (num_sms() / 2));
        int n = 0;
        e = cudaOccupancyMaxActiveClusters(&n, conv3x3_halo_kernel<BN, HT, true, STATS>, &cfg);
        if (e != cudaSuccess || n < 1) return set_error(VU_ERR_CUDA, "cudaOccupancyMaxActiveClusters(halo): %s", cudaGetErrorString(e));
        max_clusters = n;
    }
    int pairs = (hp.num_tiles + 1) / 2;
    int cap = max_clusters;
    if (max_ctas > 0 && max_ctas / 2 < cap) cap = max_ctas / 2 > 0 ? max_ctas / 2 : 1;
    if (pairs > cap) pairs = cap;
    cfg.gridDim = dim3(2 * pairs);
    cudaError_t e = cudaLaunchKernelEx(&cfg, conv3x3_halo_kernel<BN, HT, true, STATS>, xm, wm, hp);
    if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "conv3x3_halo pair launch: %s", cudaGetErrorString(e));
    return 1;
}

int conv3x3_halo_launch(const vu_conv_args* a, cudaStream_t stream) {
    const int up = a->out_up > 0 ? a->out_up : 1;
    if (!(a->KH == 3 && a->KW == 3 && a->stride == 1 && a->pad == 1 && up == 1)) return 0;
    if (!(a->Cout == 64 || a->Cout == 128)) return 0;
    if (a->tap_mask != 0 && a->tap_mask != 0x1FF) return 0;      // tap subsets run on the per-tap kernel
    static const int disabled = getenv("VU_NO_HALO") ? atoi(getenv("VU_NO_HALO")) : 0;
    if (disabled) return 0;
    if (a->W < 96) return 0;   // narrow maps: the 128-pixel tile row would be mostly padding
    // rows per tile: 4 for Cout 64 (TMEM 2 x 4 x 64 columns) unless that pads the height noticeably more than 2
    int ht = 2;
    if (a->Cout == 64) {
        const int pad4 = (a->H + 3) / 4 * 4, pad2 = (a->H + 1) / 2 * 2;
        ht = (pad4 * 10 <= pad2 * 11) ? 4 : 2;
        static const int force_ht = getenv("VU_HALO_HT") ? atoi(getenv("VU_HALO_HT")) : 0;
        if (force_ht == 2 || force_ht == 4) ht = force_ht;
        if (a->stat_sums != nullptr) ht = 2;         // the statistics flavour (shared-memory budget)
    }
    HaloParams hp{};
    hp.n_img = a->N; hp.h = a->H; hp.w = a->W;
    hp.tiles_w = (a->W + kHaloWT - 1) / kHaloWT;
    hp.tiles_h = (a->H + ht - 1) / ht;
    hp.num_tiles = hp.tiles_w * hp.tiles_h * a->N;
    hp.kblocks = (a->Cin + 63) / 64;
    hp.last_ksteps = ((a->Cin - (hp.kblocks - 1) * 64) + 15) / 16;
    hp.scale = a->scale; hp.bias = a->bias;
    hp.residual = static_cast<const uint16_t*>(a->residual); hp.res_ctot = a->res_ctot; hp.res_coff = a->res_coff;
    if (a->fuse_out && !a->fuse_w) return set_error(VU_ERR_INVALID, "conv: fuse_out without fuse_w");
    hp.y = static_cast<uint16_t*>(a->y); hp.y_ctot = a->y_ctot; hp.y_coff = a->y_coff;
    hp.relu = a->relu; hp.dtype = a->dtype;
    hp.fuse_w = a->fuse_w; hp.fuse_bias = a->fuse_bias; hp.fuse_out = a->fuse_out;
    hp.pos_bias = a->pos_bias;
    hp.stat_sums = a->stat_sums;

    // CTA pairs (cta_group::2) for the tap-reuse kernel as well; VU_HALO_PAIR=0 selects single CTAs (A/B runs)
    static const int pair_mode = getenv("VU_HALO_PAIR") ? atoi(getenv("VU_HALO_PAIR")) : 1;
    const bool pair = pair_mode != 0 && hp.num_tiles >= 2;

    CUtensorMap xm, wm;
    {
        const uint16_t* xb = static_cast<const uint16_t*>(a->x) + a->x_coff;
        const uint64_t pix = (uint64_t)a->x_ctot * 2;
        const uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->N};
        const uint64_t str[3] = {pix, pix * a->W, pix * a->W * a->H};
        const uint32_t box[4] = {64u, (uint32_t)kHaloPitch, 1u, 1u};
        int rc = encode_map(&xm, a->dtype, 4, xb, dims, str, box);
        if (rc) return rc;
    }
    {
        const uint64_t dims[3] = {(uint64_t)a->Cin, 9u, (uint64_t)a->Cout};
        const uint64_t str[2] = {(uint64_t)a->Cin * 2, (uint64_t)a->Cin * 2 * 9};
        const uint32_t wbox[3] = {64u, 1u, (uint32_t)(pair ? a->Cout / 2 : a->Cout)};
        int rc = encode_map(&wm, a->dtype, 3, a->w, dims, str, wbox);
        if (rc) return rc;
    }
    if (hp.stat_sums != nullptr) {                  // training flavour (ht was forced to 2 above)
        if (pair) return a->Cout == 64 ? launch_halo_pair<64, 2, true>(xm, wm, hp, a->max_ctas, stream)
                                       : launch_halo_pair<128, 2, true>(xm, wm, hp, a->max_ctas, stream);
        return a->Cout == 64 ? launch_halo<64, 2, true>(xm, wm, hp, a->max_ctas, stream)
                             : launch_halo<128, 2, true>(xm, wm, hp, a->max_ctas, stream);
    }
    if (pair) {
        if (a->Cout == 64) {
            if (ht == 4) return launch_halo_pair<64, 4, false>(xm, wm, hp, a->max_ctas, stream);
            return launch_halo_pair<64, 2, false>(xm, wm, hp, a->max_ctas, stream);
        }
        return launch_halo_pair<128, 2, false>(xm, wm, hp, a->max_ctas, stream);
    }
    if (a->Cout == 64) {
        if (ht == 4) return launch_halo<64, 4, false>(xm, wm, hp, a->max_ctas, stream);
        return launch_halo<64, 2, false>(xm, wm, hp, a->max_ctas, stream);
    }
    return launch_halo<128, 2, false>(xm, wm, hp, a->max_ctas, stream);
}

}  // namespace vu
