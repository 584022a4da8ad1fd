// 3x3 s1 p1 convolution on NARROW maps (W <= 64: ResNet layers 2-4 and the two low-resolution decoder blocks) - the tap
// reuse of vu_conv_halo.cu generalised with a flat padded pixel index.
//
// Why: on these layers the implicit GEMM of vu_conv_igemm.cu is bound by L2->SM traffic, not by the tensor pipe - every
// input pixel crosses the crossbar once per tap (ncu r2d: 226 MB for a 17 MB input, 5.9 TB/s, tensor pipe 24-35 %), and at
// the training batch (16 images) a layer is only 1-4 tiles per CTA. vu_conv_halo.cu removes the 9x for W >= 96 with
// tiles of 128-pixel rows; a 64/32/16-pixel row cannot fill the 128 lanes of an MMA.
//
// Here a tile is HT whole image rows staged with their halo as (HT+2) x P pixels, P = W + 2 (TMA zero fill = the
// convolution's padding, left and right columns included). Output positions are numbered FLAT over the padded rows,
// f = j * P + c, and an M = 128 MMA covers 128 CONSECUTIVE flat positions: for tap (r, s) its A operand is the same
// shared-memory tile with the descriptor start advanced by (f0 + r * P + s) rows of 128 B (the 128-byte swizzle is a function
// of the absolute address, the property vu_conv_halo.cu relies on). Positions with c >= W are junk columns between the image
// rows - 2 of every P MMA lanes - and are dropped by the epilogue; junk rows past the staged tile only feed junk positions.
// L2->SM traffic per MMA falls from 9 pixel loads to (HT+2)/HT.
//
//   GEMM view per work item: MT accumulators [128 flat positions x 128 channels] per CTA, CTA pairs (cta_group::2, M = 256),
//   K = 9 taps x Cin; work item = (pair of tiles, 128-channel N tile).  MT = 2: 256 positions per CTA (HT = 256 / P rows);
//   MT = 1: 128 positions (HT = 128 / P), chosen when the larger tile would leave the last wave mostly empty.
//   Warps: 0 = A producer (one TMA box per tile row), 1 = TMEM allocator + MMA issuer, 2 = B producer (one filter tap of
//   64 rows per stage: the pair shares a tap), 3..10 = epilogue (MT = 2: one M tile per warp group; MT = 1: column halves).
#include "vu_conv_common.cuh"
#include <cstdlib>

namespace vu {

constexpr int kFlatBN = 128;
constexpr int kFlatEpiWarps = 8;
constexpr int kFlatThreads = 96 + 32 * kFlatEpiWarps;
constexpr int kFlatMaxPitch = 66;
constexpr int kFlatAStages = 2;
constexpr int kFlatBBytes = (kFlatBN / 2) * 128;
constexpr int kFlatBStages = 8;
constexpr int kFlatMaxCout = 1024;

struct FlatParams {
    int n_img, h, w, pitch, ht, tiles_h, num_tiles, n_tiles, num_items;
    int kblocks, last_ksteps, cout;
    const float* scale; const float* bias;
    const uint16_t* residual; int res_ctot, res_coff;
    uint16_t* y; int y_ctot, y_coff;
    int relu, dtype;
    double* stat_sums;
};

template <int MT> struct FlatCfg {
    static constexpr int kARows = (MT * 128 + 2 * kFlatMaxPitch + 2 + 7) / 8 * 8;          // rows of 128 B per stage
    static constexpr int kAStride = (kARows * 128 + 1023) / 1024 * 1024;
    static constexpr int kTmemCols = 2 * MT * kFlatBN;
    static constexpr int kBarBytes = 256;
    static constexpr int kStatBytes = kFlatEpiWarps * 2 * kEpiStatMaxCout * 4 + kFlatEpiWarps * kEpiScratchBytes;
    static constexpr int kSmemBytes = 1024 + kFlatAStages * kAStride + kFlatBStages * kFlatBBytes + kBarBytes +
                                      2 * kFlatMaxCout * 4 + kStatBytes;
    static_assert(kTmemCols <= 512, "TMEM budget");
    static_assert(kSmemBytes <= 232448, "smem budget");
};

template <int MT, bool STATS>
__global__ void __launch_bounds__(kFlatThreads, 1)
conv3x3_flat_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w, const FlatParams p) {
    using Cfg = FlatCfg<MT>;
    constexpr int BN = kFlatBN;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
    const uint32_t a_base = smem_base;
    const uint32_t b_base = smem_base + kFlatAStages * Cfg::kAStride;
    constexpr int kBarOff = kFlatAStages * Cfg::kAStride + kFlatBStages * kFlatBBytes;
    const uint32_t bar_base = smem_base + kBarOff;
    auto a_full = [&](int s) { return bar_base + 8u * s; };
    auto a_empty = [&](int s) { return bar_base + 8u * (kFlatAStages + s); };
    auto b_full = [&](int s) { return bar_base + 8u * (2 * kFlatAStages + s); };
    auto b_empty = [&](int s) { return bar_base + 8u * (2 * kFlatAStages + kFlatBStages + s); };
    auto t_full = [&](int a) { return bar_base + 8u * (2 * kFlatAStages + 2 * kFlatBStages + a); };
    auto t_empty = [&](int a) { return bar_base + 8u * (2 * kFlatAStages + 2 * kFlatBStages + 2 + a); };
    constexpr int kSlotIdx = 2 * kFlatAStages + 2 * kFlatBStages + 4;
    static_assert(8 * (kSlotIdx + 1) <= Cfg::kBarBytes, "barrier area");
    const uint32_t tmem_slot = bar_base + 8u * kSlotIdx;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_al + kBarOff + 8 * kSlotIdx);
    float* s_scale = reinterpret_cast<float*>(smem_al + kBarOff + Cfg::kBarBytes);
    float* s_bias = s_scale + kFlatMaxCout;
    float* s_stat = s_bias + kFlatMaxCout;                 // STATS: [8 epilogue warps][2][kEpiStatMaxCout] fp32 partials
    uint8_t* s_scratch = reinterpret_cast<uint8_t*>(s_stat + kFlatEpiWarps * 2 * kEpiStatMaxCout);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    // work item = (pair of tiles (2t, 2t+1), N tile); full barriers live in the leader (rank 0) and take both producers,
    // empty / accumulator-full barriers are per CTA (multicast commit), accumulator-empty lives in the leader
    const uint32_t rank = cluster_ctarank();
    const int item0 = (int)(blockIdx.x >> 1), item_step = (int)(gridDim.x >> 1);
    const int P = p.pitch;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_w);
        for (int s = 0; s < kFlatAStages; ++s) { mbar_init(a_full(s), 2); mbar_init(a_empty(s), 1); }
        for (int s = 0; s < kFlatBStages; ++s) { mbar_init(b_full(s), 2); mbar_init(b_empty(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(t_full(a), 1); mbar_init(t_empty(a), 2 * 32 * kFlatEpiWarps); }
        fence_barrier_init();
    }
    if (warp == 1) { tmem_alloc_pair<Cfg::kTmemCols>(tmem_slot); tmem_relinquish_pair(); }
    for (int c = threadIdx.x; c < p.cout; c += kFlatThreads) {
        s_scale[c] = p.scale ? p.scale[c] : 1.0f;
        s_bias[c] = p.bias ? p.bias[c] : 0.0f;
    }
    if constexpr (STATS)
        for (int c = threadIdx.x; c < kFlatEpiWarps * 2 * kEpiStatMaxCout; c += kFlatThreads) s_stat[c] = 0.0f;
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        // ===================== A producer: the (HT+2) x P halo'd rows of this CTA's tile, one TMA box per row ==========
        if (lane == 0) {
            uint32_t it = 0;
            const uint32_t leader_a_full0 = mapa_cluster(a_full(0), 0);
            const uint32_t row_bytes = (uint32_t)P * 128u;
            for (int item = item0; item < p.num_items; item += item_step) {
                const int tile = 2 * (item / p.n_tiles) + (int)rank;          // may be one past the end (loads are zero fill)
                const int hb = tile % p.tiles_h, nb = tile / p.tiles_h;
                for (int kb = 0; kb < p.kblocks; ++kb, ++it) {
                    const int s = it % kFlatAStages;
                    mbar_wait(a_empty(s), ((it / kFlatAStages) & 1) ^ 1);
                    if (rank == 0) mbar_arrive_expect_tx(a_full(s), 2u * (uint32_t)(p.ht + 2) * row_bytes);
                    else mbar_arrive_cluster(leader_a_full0 + 8u * s);
                    for (int i = 0; i < p.ht + 2; ++i)
                        tma_load_4d_pair(a_base + s * Cfg::kAStride + i * row_bytes, &map_x, a_full(s), kb * 64, -1,
                                         hb * p.ht - 1 + i, nb);
                }
            }
        }
    } else if (warp == 2) {
        // ===================== B producer: one tap of the N tile per stage, half of its 128 filter rows per CTA ==========
        if (lane == 0) {
            uint32_t it = 0;
            const uint32_t leader_b_full0 = mapa_cluster(b_full(0), 0);
            for (int item = item0; item < p.num_items; item += item_step) {
                const int n0 = (item % p.n_tiles) * BN;
                for (int kb = 0; kb < p.kblocks; ++kb) {
                    for (int tap = 0; tap < 9; ++tap, ++it) {
                        const int s = it % kFlatBStages;
                        mbar_wait(b_empty(s), ((it / kFlatBStages) & 1) ^ 1);
                        if (rank == 0) mbar_arrive_expect_tx(b_full(s), 2 * kFlatBBytes);
                        else mbar_arrive_cluster(leader_b_full0 + 8u * s);
                        tma_load_3d_pair(b_base + s * kFlatBBytes, &map_w, b_full(s), kb * 64, tap, n0 + (int)rank * (BN / 2));
                    }
                }
            }
        }
    } else if (warp == 1 && rank == 0) {
        // ===================== MMA issuer (leader CTA): M = 256 over the pair, MT accumulators per CTA ==================
        const uint32_t idesc = umma_idesc_f16(256, BN, p.dtype == 0 ? 1u : 0u);
        uint32_t ita = 0, itb = 0, tcount = 0;
        for (int item = item0; item < p.num_items; item += item_step, ++tcount) {
            const int acc = tcount & 1;
            mbar_wait(t_empty(acc), ((tcount >> 1) & 1) ^ 1);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc * (MT * BN);
            for (int kb = 0; kb < p.kblocks; ++kb, ++ita) {
                const int sa = ita % kFlatAStages;
                mbar_wait(a_full(sa), (ita / kFlatAStages) & 1);
                tc_fence_after();
                const uint64_t a0 = umma_desc_k_sw128(a_base + sa * Cfg::kAStride);
                const int nk = (kb == p.kblocks - 1) ? p.last_ksteps : 4;
#pragma unroll
                for (int tap = 0; tap < 9; ++tap, ++itb) {
                    const int sb = itb % kFlatBStages;
                    mbar_wait(b_full(sb), (itb / kFlatBStages) & 1);
                    tc_fence_after();
                    const uint64_t bdesc = umma_desc_k_sw128(b_base + sb * kFlatBBytes);
                    const int r = tap / 3, s = tap - 3 * r;
                    const uint32_t shift = (uint32_t)(r * P + s) * 8u;        // rows of 128 B in 16-byte units
                    if (elect_one()) {
                        for (int k = 0; k < nk; ++k)
#pragma unroll
                            for (int m = 0; m < MT; ++m)
                                umma_f16_ss_pair(tmem_d + m * BN, a0 + shift + (uint32_t)(m * 128 * 8 + 2 * k), bdesc + 2u * k, idesc,
                                                 (kb | tap | k) != 0);
                        umma_commit_pair(b_empty(sb), 3);
                        if (tap == 8) umma_commit_pair(a_empty(sa), 3);
                    }
                    __syncwarp();
                }
            }
            if (elect_one()) umma_commit_pair(t_full(acc), 3);
            __syncwarp();
        }
    }
    else if (warp >= 3) {
        // ===================== epilogue: warps 3..10, two per TMEM lane quarter ==========================================
        const int q = warp & 3;
        const int half = (warp - 3) >> 2;
        const int mt = (MT == 2) ? half : 0;                   // MT = 2: one M tile per warp group; MT = 1: column halves
        const int c_begin = (MT == 2) ? 0 : 2 * half, c_end = (MT == 2) ? BN / 32 : 2 * half + 2;
        const int f = mt * 128 + q * 32 + lane;                // flat position in the padded tile
        const int j = f / P, c = f - j * P;
        const bool has_res = p.residual != nullptr, relu = p.relu != 0;
        const bool affine = p.scale != nullptr || p.bias != nullptr;
        const int dtype = p.dtype;
        uint8_t* scratch = s_scratch + (warp - 3) * kEpiScratchBytes;
        float* my_sum = s_stat + ((warp - 3) * 2) * kEpiStatMaxCout;
        float* my_sq = my_sum + kEpiStatMaxCout;
        uint32_t tcount = 0;
        const uint32_t leader_t_empty0 = mapa_cluster(t_empty(0), 0);
        for (int item = item0; item < p.num_items; item += item_step, ++tcount) {
            const int tile = 2 * (item / p.n_tiles) + (int)rank;
            const int n0 = (item % p.n_tiles) * BN;
            const int acc = tcount & 1;
            const int hb = tile % p.tiles_h, nb = tile / p.tiles_h;
            const int ho = hb * p.ht + j;
            const bool valid = tile < p.num_tiles && j < p.ht && c < p.w && ho < p.h;
            const size_t pix = (static_cast<size_t>(nb) * p.h + ho) * p.w + c;
            mbar_wait(t_full(acc), (tcount >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * (MT * BN) + mt * BN;
#pragma unroll 1
            for (int cc = c_begin; cc < c_end; ++cc) {
                const int ch0 = n0 + cc * 32;
                if (ch0 >= p.cout) break;                          // partial last N tile (Cout % 32 == 0)
                uint32_t r[32];
                tmem_ld_32x32b_x32(taddr + cc * 32, r);
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
                    for (int i = 0; i < 4; ++i) op[i] = make_uint4(outw[4 * i], outw[4 * i + 1], outw[4 * i + 2], outw[4 * i + 3]);
                }
                if constexpr (STATS) {
                    if (p.stat_sums != nullptr)                  // statistics of the values as stored (junk positions count as zeros)
                        epi_stats_chunk(outw, valid, lane, dtype, scratch, my_sum + ch0, my_sq + ch0);
                }
            }
            tc_fence_before();
            mbar_arrive_cluster(leader_t_empty0 + 8u * acc);
        }
    }

    tc_fence_before();
    cluster_sync_all();
    if constexpr (STATS)
        if (p.stat_sums != nullptr) epi_stats_flush(s_stat, p.stat_sums, p.cout, threadIdx.x, kFlatThreads, kEpiStatMaxCout, kFlatEpiWarps);
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
    }
}

template <int MT, bool STATS>
static int launch_flat(const CUtensorMap& xm, const CUtensorMap& wm, const FlatParams& fp, int max_ctas, cudaStream_t stream) {
    using Cfg = FlatCfg<MT>;
    static int max_clusters = -1;
    cudaLaunchConfig_t cfg{};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.blockDim = dim3(kFlatThreads); cfg.dynamicSmemBytes = Cfg::kSmemBytes; cfg.stream = stream;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (max_clusters < 0) {
        cudaError_t e = cudaFuncSetAttribute(conv3x3_flat_kernel<MT, STATS>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "cudaFuncSetAttribute(flat): %s", cudaGetErrorString(e));
        cfg.gridDim = dim3(2 * (num_sms() / 2));
        int n = 0;
        e = cudaOccupancyMaxActiveClusters(&n, conv3x3_flat_kernel<MT, STATS>, &cfg);
        if (e != cudaSuccess || n < 1) return set_error(VU_ERR_CUDA, "cudaOccupancyMaxActiveClusters(flat): %s", cudaGetErrorString(e));
        max_clusters = n;
    }
    int pairs = fp.num_items;
    int cap = max_clusters;
    if (max_ctas > 0 && max_ctas / 2 < cap) cap = max_ctas / 2 > 0 ? max_ctas / 2 : 1;
    if (pairs > cap) pairs = cap;
    cfg.gridDim = dim3(2 * pairs);
    cudaError_t e = cudaLaunchKernelEx(&cfg, conv3x3_flat_kernel<MT, STATS>, xm, wm, fp);
    if (e != cudaSuccess) return set_error(VU_ERR_CUDA, "conv3x3_flat launch: %s", cudaGetErrorString(e));
    return 1;
}

// returns 1 when it took the launch, 0 when the shape is not its case, < 0 on error
int conv3x3_flat_launch(const vu_conv_args* a, cudaStream_t stream) {
    const int up = a->out_up > 0 ? a->out_up : 1;
    if (!(a->KH == 3 && a->KW == 3 && a->stride == 1 && a->pad == 1 && up == 1)) return 0;
    if (a->tap_mask != 0 && a->tap_mask != 0x1FF) return 0;
    if (a->pos_bias != nullptr || a->fuse_out != nullptr || a->y == nullptr) return 0;
    const char* mode_env = getenv("VU_FLAT");                                   // read per call: tests switch it
    const int mode = mode_env ? atoi(mode_env) : 1;                             // 0 = off, 1 = auto, 2 = every shape that fits
    if (mode == 0) return 0;
    if (a->W + 2 > kFlatMaxPitch || a->W < 8 || a->H < 4) return 0;
    if (a->Cout % 32 || a->Cout > kFlatMaxCout || a->Cout < 64) return 0;
    if (a->stat_sums != nullptr && a->Cout > kEpiStatMaxCout) return 0;
    // auto: Cout <= 128 only. Measured alone (profiles/r2zh_flat_vs_pair.log, N = 16, f16): 128 -> 128 @64^2 611 vs 510 TFLOP/s
    // for the CTA-pair implicit GEMM (an A element feeds only 128 MACs per crossbar trip there), but with 256-column tiles
    // the implicit GEMM is already at 1 100 - 1 350 TFLOP/s and this kernel's 128-column tiles re-read the filters twice as
    // often (256 -> 256 @64^2: 1 034 vs 1 144; 800 -> 512 @32^2: 976 vs 1 282). VU_FLAT=2 forces it for every shape that fits.
    const size_t npix = (size_t)a->N * a->H * a->W;
    if (mode == 1 && (npix < 4096 || a->Cout > 128)) return 0;
    FlatParams fp{};
    fp.n_img = a->N; fp.h = a->H; fp.w = a->W; fp.pitch = a->W + 2;
    fp.kblocks = (a->Cin + 63) / 64;
    fp.last_ksteps = ((a->Cin - (fp.kblocks - 1) * 64) + 15) / 16;
    fp.cout = a->Cout; fp.n_tiles = (a->Cout + kFlatBN - 1) / kFlatBN;
    // tile height: MT = 2 (256 positions per CTA) or MT = 1 (128). Estimated cost = waves x MMA time of a tile
    int best_mt = 0, best_ht = 0;
    double best_cost = 1e30;
    const int pairs_cap = num_sms() / 2;
    for (int mt = 2; mt >= 2; --mt) {      // MT = 1 measured slower on every shape; kept selectable with VU_FLAT_MT=1
        int ht = (mt * 128) / fp.pitch;
        if (ht > a->H) ht = a->H;
        if (ht < 1) continue;
        const int tiles_h = (a->H + ht - 1) / ht;
        const long items = (long)((a->N * tiles_h + 1) / 2) * fp.n_tiles;
        const long waves = (items + pairs_cap - 1) / pairs_cap;
        const double cost = (double)waves * (mt + 0.15);                           // + pipeline fill / epilogue tail per tile
        if (cost < best_cost) { best_cost = cost; best_mt = mt; best_ht = ht; }
    }
    if (best_mt == 0) return 0;
    const int force_mt = getenv("VU_FLAT_MT") ? atoi(getenv("VU_FLAT_MT")) : 0;
    if (force_mt == 1 || force_mt == 2) {
        best_mt = force_mt;
        best_ht = (force_mt * 128) / fp.pitch;
        if (best_ht > a->H) best_ht = a->H;
        if (best_ht < 1) return 0;
    }
    fp.ht = best_ht;
    fp.tiles_h = (a->H + fp.ht - 1) / fp.ht;
    fp.num_tiles = fp.tiles_h * a->N;
    fp.num_items = ((fp.num_tiles + 1) / 2) * fp.n_tiles;
    fp.scale = a->scale; fp.bias = a->bias;
    fp.residual = static_cast<const uint16_t*>(a->residual); fp.res_ctot = a->res_ctot; fp.res_coff = a->res_coff;
    fp.y = static_cast<uint16_t*>(a->y); fp.y_ctot = a->y_ctot; fp.y_coff = a->y_coff;
    fp.relu = a->relu; fp.dtype = a->dtype;
    fp.stat_sums = a->stat_sums;

    CUtensorMap xm, wm;
    {
        const uint16_t* xb = static_cast<const uint16_t*>(a->x) + a->x_coff;
        const uint64_t pix = (uint64_t)a->x_ctot * 2;
        const uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->N};
        const uint64_t str[3] = {pix, pix * a->W, pix * a->W * a->H};
        const uint32_t box[4] = {64u, (uint32_t)fp.pitch, 1u, 1u};
        int rc = encode_map(&xm, a->dtype, 4, xb, dims, str, box);
        if (rc) return rc;
    }
    {
        const uint64_t dims[3] = {(uint64_t)a->Cin, 9u, (uint64_t)a->Cout};
        const uint64_t str[2] = {(uint64_t)a->Cin * 2, (uint64_t)a->Cin * 2 * 9};
        const uint32_t wbox[3] = {64u, 1u, (uint32_t)(kFlatBN / 2)};
        int rc = encode_map(&wm, a->dtype, 3, a->w, dims, str, wbox);
        if (rc) return rc;
    }
    if (fp.stat_sums != nullptr)
        return best_mt == 2 ? launch_flat<2, true>(xm, wm, fp, a->max_ctas, stream) : launch_flat<1, true>(xm, wm, fp, a->max_ctas, stream);
    return best_mt == 2 ? launch_flat<2, false>(xm, wm, fp, a->max_ctas, stream) : launch_flat<1, false>(xm, wm, fp, a->max_ctas, stream);
}

}  // namespace vu
