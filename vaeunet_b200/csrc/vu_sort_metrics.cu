// Sort-based uncertainty metrics on the device (reference utils/uncertainty_metrics.py:175-240 sparsification
// curve, :270-283 uncertainty-vs-error AUROC / AUPRC, which run numpy argsort + sklearn on the host):
//   * a stable LSD radix sort of (float key, 32-bit payload) pairs, DESCENDING by key - 8 passes of 4 bits: per-block digit
//     histograms, one exclusive scan over the digit-major histogram, and a stable scatter whose in-block ranks come from
//     per-thread digit counters prefix-summed across the block (blocked arrangement: thread t owns 8 consecutive elements);
//   * per-pixel NLL errors + their summary statistics, sums of the sorted errors between the curve's cut points, and sums
//     under a host-drawn keep-mask (the random baseline uses numpy's RNG stream by the reference's definition);
//   * the binary-classification curve of sklearn's _binary_clf_curve (cumulative TP / FP at the LAST element of every group
//     of tied scores) by one prefix scan + stream compaction, integrated to AUROC (trapezoid) and average precision (step).
// All kernels are memory-bound streaming passes; accumulators that cross blocks are fp64.
#include "vu_internal.h"
#include "vu_half.cuh"

namespace vu {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 8;
constexpr int kSortTile = kSortThreads * kSortItems;      // 2048 elements per block
constexpr int kRadixBits = 4;
constexpr int kRadix = 1 << kRadixBits;

// float -> unsigned key whose ASCENDING order is the DESCENDING order of the floats (-0 < +0 like the total order; NaNs last)
__device__ __forceinline__ uint32_t desc_key(float f) {
    uint32_t u = __float_as_uint(f);
    u ^= (u >> 31) ? 0xFFFFFFFFu : 0x80000000u;           // ascending-order key
    return ~u;
}

__global__ void __launch_bounds__(256)
make_keys_kernel(const float* __restrict__ score, size_t n, uint32_t* __restrict__ keys) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        keys[i] = desc_key(score[i]);
}

__global__ void __launch_bounds__(kSortThreads)
radix_hist_kernel(const uint32_t* __restrict__ keys, size_t n, int shift, uint32_t* __restrict__ hist, int num_blocks) {
    __shared__ uint32_t h[kRadix];
    if (threadIdx.x < kRadix) h[threadIdx.x] = 0;
    __syncthreads();
    const size_t base = (size_t)blockIdx.x * kSortTile + (size_t)threadIdx.x * kSortItems;
#pragma unroll
    for (int j = 0; j < kSortItems; ++j)
        if (base + j < n) atomicAdd(&h[(keys[base + j] >> shift) & (kRadix - 1)], 1u);
    __syncthreads();
    if (threadIdx.x < kRadix) hist[(size_t)threadIdx.x * num_blocks + blockIdx.x] = h[threadIdx.x];
}

// exclusive prefix sum of a uint32 array in place, one block (arrays of up to a few hundred thousand entries)
__global__ void __launch_bounds__(1024)
scan_u32_single_block_kernel(uint32_t* __restrict__ a, size_t n) {
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (size_t base = 0; base < n; base += 1024) {
        const size_t i = base + threadIdx.x;
        const uint32_t v = i < n ? a[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = warp_tot[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            warp_tot[lane] = w;                             // inclusive totals of the warps
        }
        __syncthreads();
        const uint32_t carry = carry_s;
        const uint32_t excl = carry + (warp ? warp_tot[warp - 1] : 0u) + x - v;
        if (i < n) a[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + warp_tot[31];
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kSortThreads)
radix_scatter_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals, size_t n, int shift,
                     const uint32_t* __restrict__ hist_scanned, int num_blocks, uint32_t* __restrict__ keys_out,
                     uint32_t* __restrict__ vals_out) {
    __shared__ uint16_t cnt[kRadix][kSortThreads];          // per (digit, thread): elements of that digit owned by the thread
    __shared__ uint32_t base_d[kRadix];
    const int t = threadIdx.x;
    const size_t base = (size_t)blockIdx.x * kSortTile + (size_t)t * kSortItems;
    uint32_t k[kSortItems], v[kSortItems];
#pragma unroll
    for (int d = 0; d < kRadix; ++d) cnt[d][t] = 0;
#pragma unroll
    for (int j = 0; j < kSortItems; ++j)
        if (base + j < n) {
            k[j] = keys[base + j]; v[j] = vals[base + j];
            cnt[(k[j] >> shift) & (kRadix - 1)][t] += 1;    // this thread's own column: no atomics
        }
    if (t < kRadix) base_d[t] = hist_scanned[(size_t)t * num_blocks + blockIdx.x];
    __syncthreads();
    // exclusive scan over the threads of every digit row: warp w takes digits w and w + 8
    const int lane = t & 31, warp = t >> 5;
    for (int d = warp; d < kRadix; d += kSortThreads / 32) {
        uint32_t carry = 0;
        for (int c0 = 0; c0 < kSortThreads; c0 += 32) {
            const uint32_t val = cnt[d][c0 + lane];
            uint32_t x = val;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
                if (lane >= o) x += y;
            }
            cnt[d][c0 + lane] = (uint16_t)(carry + x - val);
            carry += __shfl_sync(0xffffffffu, x, 31);
        }
    }
    __syncthreads();
    // stable placement: elements in (thread, item) order = input order
#pragma unroll
    for (int j = 0; j < kSortItems; ++j)
        if (base + j < n) {
            const int d = (k[j] >> shift) & (kRadix - 1);
            const uint32_t pos = base_d[d] + cnt[d][t];
            cnt[d][t] += 1;
            keys_out[pos] = k[j];
            vals_out[pos] = v[j];
        }
}

// ------------------------------------------------------------------------------------------- sparsification pieces
// err = -(g log(p + eps) + (1 - g) log(1 - p + eps)); stats[0] += sum err (fp64), stats[1] = min, stats[2] = max (as ordered
// uint bit patterns via atomics on floats >= 0), stats[3] += number of NaNs
__global__ void __launch_bounds__(256)
pixel_nll_kernel(const float* __restrict__ p, const float* __restrict__ g, size_t n, float eps, float* __restrict__ err,
                 double* __restrict__ sum_nan, unsigned int* __restrict__ minmax) {
    float s = 0.f, mn = INFINITY, mx = -INFINITY;
    unsigned int nans = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float pi = p[i], gi = g[i];
        const float e = -(gi * logf(pi + eps) + (1.f - gi) * logf(1.f - pi + eps));
        err[i] = e;
        if (e != e) ++nans; else { s += e; mn = fminf(mn, e); mx = fmaxf(mx, e); }
    }
    s = warp_sum(s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        nans += __shfl_xor_sync(0xffffffffu, nans, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&sum_nan[0], (double)s);
        if (nans) atomicAdd(&sum_nan[1], (double)nans);
        // order-preserving unsigned encoding of any float (errors can be slightly negative when p + eps > 1)
        auto enc = [](float f) { uint32_t u = __float_as_uint(f); return u ^ ((u >> 31) ? 0xFFFFFFFFu : 0x80000000u); };
        if (mn <= mx) { atomicMin(&minmax[0], enc(mn)); atomicMax(&minmax[1], enc(mx)); }
    }
}

// sums[b] += sum of x[i] (float bit patterns) for cuts[b] <= i < cuts[b + 1], b < nb (cuts ascending, cuts[nb] = n)
__global__ void __launch_bounds__(256)
sorted_bin_sums_kernel(const uint32_t* __restrict__ x, size_t n, const long long* __restrict__ cuts, int nb,
                       double* __restrict__ sums) {
    __shared__ long long sc[65];
    for (int i = threadIdx.x; i <= nb; i += blockDim.x) sc[i] = cuts[i];
    __syncthreads();
    int bin = -1;
    double acc = 0.0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        int b = bin < 0 ? 0 : bin;
        while (b + 1 < nb && (long long)i >= sc[b + 1]) ++b;       // indices only grow along a thread's stride
        if (b != bin) {
            if (bin >= 0 && acc != 0.0) atomicAdd(&sums[bin], acc);
            bin = b; acc = 0.0;
        }
        acc += (double)__uint_as_float(x[i]);
    }
    if (bin >= 0 && acc != 0.0) atomicAdd(&sums[bin], acc);
}

// out[0] += sum of x[i] over the set bits of a packed keep-mask (bit i & 7 of byte i >> 3)
__global__ void __launch_bounds__(256)
masked_sum_kernel(const float* __restrict__ x, const uint8_t* __restrict__ bits, size_t n, double* __restrict__ out) {
    double acc = 0.0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        if ((bits[i >> 3] >> (i & 7)) & 1) acc += (double)x[i];
    acc = warp_sum_d(acc);
    if ((threadIdx.x & 31) == 0 && acc != 0.0) atomicAdd(out, acc);
}

// ------------------------------------------------------------------------------------------- binary-classification curve
// label[i] = ((p[i] > thr) != (g[i] > 0.5)) as 0 / 1 (uncertainty_metrics.py:271-275)
__global__ void __launch_bounds__(256)
error_labels_kernel(const float* __restrict__ p, const float* __restrict__ g, size_t n, float thr, uint32_t* __restrict__ lab) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        lab[i] = ((p[i] > thr) != (g[i] > 0.5f)) ? 1u : 0u;
}

// Pass 1: per block of 2048 sorted elements, (number of positives, number of group ends). A group end is the LAST element of
// a run of equal keys (sklearn's distinct_value_indices + the final element).
__global__ void __launch_bounds__(kSortThreads)
curve_block_counts_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ lab, size_t n,
                          uint32_t* __restrict__ blk_pos, uint32_t* __restrict__ blk_ends) {
    const size_t base = (size_t)blockIdx.x * kSortTile + (size_t)threadIdx.x * kSortItems;
    uint32_t pos = 0, ends = 0;
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const size_t i = base + j;
        if (i < n) {
            pos += lab[i];
            ends += (i + 1 == n || keys[i] != keys[i + 1]) ? 1u : 0u;
        }
    }
    pos = (uint32_t)__reduce_add_sync(0xffffffffu, pos);
    ends = (uint32_t)__reduce_add_sync(0xffffffffu, ends);
    __shared__ uint32_t sp[8], se[8];
    if ((threadIdx.x & 31) == 0) { sp[threadIdx.x >> 5] = pos; se[threadIdx.x >> 5] = ends; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t a = 0, b = 0;
        for (int w = 0; w < 8; ++w) { a += sp[w]; b += se[w]; }
        blk_pos[blockIdx.x] = a; blk_ends[blockIdx.x] = b;
    }
}

// Pass 2 (after the exclusive scans of the two block arrays): curve points (tps, fps) at every group end, compacted in order
__global__ void __launch_bounds__(kSortThreads)
curve_points_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ lab, size_t n,
                    const uint32_t* __restrict__ blk_pos_x, const uint32_t* __restrict__ blk_ends_x,
                    uint2* __restrict__ pts) {
    __shared__ uint32_t s_pos[kSortThreads], s_end[kSortThreads];
    const int t = threadIdx.x;
    const size_t base = (size_t)blockIdx.x * kSortTile + (size_t)t * kSortItems;
    uint32_t l[kSortItems], e[kSortItems], pos = 0, ends = 0;
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const size_t i = base + j;
        l[j] = 0; e[j] = 0;
        if (i < n) {
            l[j] = lab[i];
            e[j] = (i + 1 == n || keys[i] != keys[i + 1]) ? 1u : 0u;
        }
        pos += l[j]; ends += e[j];
    }
    s_pos[t] = pos; s_end[t] = ends;
    __syncthreads();
    // exclusive scan over the 256 threads (two arrays), Hillis-Steele in shared memory
    for (int o = 1; o < kSortThreads; o <<= 1) {
        const uint32_t a = t >= o ? s_pos[t - o] : 0u, b = t >= o ? s_end[t - o] : 0u;
        __syncthreads();
        s_pos[t] += a; s_end[t] += b;
        __syncthreads();
    }
    uint32_t cp = blk_pos_x[blockIdx.x] + s_pos[t] - pos;       // positives before this thread's first element
    uint32_t ce = blk_ends_x[blockIdx.x] + s_end[t] - ends;     // group ends before it
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const size_t i = base + j;
        if (i < n) {
            cp += l[j];
            if (e[j]) { pts[ce] = make_uint2(cp, (uint32_t)(i + 1) - cp); ++ce; }
        }
    }
}

// out[0] += sum_k (fp_k - fp_{k-1}) (tp_k + tp_{k-1}) / 2   (ROC trapezoid, un-normalised)
// out[1] += sum_k (tp_k - tp_{k-1}) tp_k / (tp_k + fp_k)      (average precision, un-normalised)
// out[2] = positives, out[3] = negatives. The number of curve points lives on the device (*npts_ptr).
__global__ void __launch_bounds__(256)
curve_integrate_kernel(const uint2* __restrict__ pts, const uint32_t* __restrict__ npts_ptr, const uint32_t* __restrict__ pos_ptr,
                       size_t n, double* __restrict__ out) {
    const size_t npts = *npts_ptr;
    double a = 0.0, b = 0.0;
    for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < npts; k += (size_t)gridDim.x * blockDim.x) {
        const uint2 c = pts[k];
        const uint2 p = k ? pts[k - 1] : make_uint2(0u, 0u);
        a += (double)(c.y - p.y) * ((double)c.x + (double)p.x) * 0.5;
        b += (double)(c.x - p.x) * ((double)c.x / ((double)c.x + (double)c.y));
    }
    a = warp_sum_d(a); b = warp_sum_d(b);
    if ((threadIdx.x & 31) == 0) {
        if (a != 0.0) atomicAdd(&out[0], a);
        if (b != 0.0) atomicAdd(&out[1], b);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { out[2] = (double)*pos_ptr; out[3] = (double)n - (double)*pos_ptr; }
}

static int sm_grid(size_t n, int per_block = 256 * 8) {
    size_t g = (n + per_block - 1) / per_block;
    const size_t cap = (size_t)num_sms() * 8;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

static size_t align256(size_t v) { return (v + 255) / 256 * 256; }

}  // namespace vu

using namespace vu;

extern "C" size_t vu_sort_workspace_bytes(size_t n) {
    const size_t nb = (n + kSortTile - 1) / kSortTile;
    // ping-pong key / value buffers + digit-major histogram + curve block arrays + curve points
    return 2 * align256(n * 4) + 2 * align256(n * 4) + align256((size_t)kRadix * nb * 4) + 2 * align256(nb * 4) + align256(n * 8) + 1024;
}

// Sort (score, payload) pairs by score DESCENDING (stable). keys_sorted receives the order-preserving unsigned keys (equal
// keys <=> equal scores), vals_sorted the payloads. n < 2^32.
extern "C" int vu_sort_pairs_desc_f32(const float* score, const uint32_t* vals, size_t n, uint32_t* keys_sorted,
                                      uint32_t* vals_sorted, void* workspace, size_t workspace_bytes, vu_stream_t st) {
    if (!score || !vals || !keys_sorted || !vals_sorted || !workspace) return set_error(VU_ERR_INVALID, "sort_pairs: null");
    if (n < 1 || n >= (1ull << 32)) return set_error(VU_ERR_INVALID, "sort_pairs: need 1 <= n < 2^32");
    if (workspace_bytes < vu_sort_workspace_bytes(n)) return set_error(VU_ERR_INVALID, "sort_pairs: workspace too small");
    cudaStream_t s = (cudaStream_t)st;
    const int nb = (int)((n + kSortTile - 1) / kSortTile);
    uint8_t* w = static_cast<uint8_t*>(workspace);
    uint32_t* ka = reinterpret_cast<uint32_t*>(w); w += align256(n * 4);
    uint32_t* kb = reinterpret_cast<uint32_t*>(w); w += align256(n * 4);
    uint32_t* va = reinterpret_cast<uint32_t*>(w); w += align256(n * 4);
    uint32_t* vb = reinterpret_cast<uint32_t*>(w); w += align256(n * 4);
    uint32_t* hist = reinterpret_cast<uint32_t*>(w);
    make_keys_kernel<<<sm_grid(n), 256, 0, s>>>(score, n, ka);
    const uint32_t* vin = vals;
    uint32_t *kin = ka, *kout = kb, *vout = va;
    for (int pass = 0; pass < 32 / kRadixBits; ++pass) {
        const int shift = pass * kRadixBits;
        const bool last = pass == 32 / kRadixBits - 1;
        if (last) { kout = keys_sorted; vout = vals_sorted; }
        radix_hist_kernel<<<nb, kSortThreads, 0, s>>>(kin, n, shift, hist, nb);
        scan_u32_single_block_kernel<<<1, 1024, 0, s>>>(hist, (size_t)kRadix * nb);
        radix_scatter_kernel<<<nb, kSortThreads, 0, s>>>(kin, vin, n, shift, hist, nb, kout, vout);
        // next pass reads what this one wrote
        uint32_t* knext_out = (kout == kb) ? ka : kb;
        uint32_t* vnext_out = (vout == va) ? vb : va;
        kin = kout; vin = vout; kout = knext_out; vout = vnext_out;
    }
    VU_CHECK_LAUNCH("sort_pairs");
    return 0;
}

extern "C" int vu_pixel_nll_errors(const float* p, const float* gt, size_t n, float eps, float* err, double* stats,
                                   vu_stream_t st) {
    if (!p || !gt || !err || !stats || n < 1) return set_error(VU_ERR_INVALID, "pixel_nll_errors: bad args");
    cudaStream_t s = (cudaStream_t)st;
    // stats (4 doubles): [0] sum, [1] NaN count, [2] and [3] hold the encoded min / max in their low 32 bits
    cudaMemsetAsync(stats, 0, 2 * sizeof(double), s);
    unsigned int* mm = reinterpret_cast<unsigned int*>(stats + 2);
    const unsigned int init[4] = {0xFFFFFFFFu, 0u, 0u, 0u};
    cudaMemcpyAsync(mm, init, sizeof(init), cudaMemcpyHostToDevice, s);
    pixel_nll_kernel<<<sm_grid(n), 256, 0, s>>>(p, gt, n, eps, err, stats, mm);
    VU_CHECK_LAUNCH("pixel_nll_errors");
    return 0;
}

extern "C" int vu_sorted_bin_sums(const uint32_t* vals_sorted, size_t n, const long long* cuts, int nbins, double* sums,
                                  vu_stream_t st) {
    if (!vals_sorted || !cuts || !sums || nbins < 1 || nbins > 64) return set_error(VU_ERR_INVALID, "sorted_bin_sums: bad args (nbins <= 64)");
    cudaStream_t s = (cudaStream_t)st;
    cudaMemsetAsync(sums, 0, nbins * sizeof(double), s);
    sorted_bin_sums_kernel<<<sm_grid(n), 256, 0, s>>>(vals_sorted, n, cuts, nbins, sums);
    VU_CHECK_LAUNCH("sorted_bin_sums");
    return 0;
}

extern "C" int vu_masked_sum(const float* x, const uint8_t* bits, size_t n, double* out, vu_stream_t st) {
    if (!x || !bits || !out) return set_error(VU_ERR_INVALID, "masked_sum: null");
    cudaStream_t s = (cudaStream_t)st;
    cudaMemsetAsync(out, 0, sizeof(double), s);
    masked_sum_kernel<<<sm_grid(n), 256, 0, s>>>(x, bits, n, out);
    VU_CHECK_LAUNCH("masked_sum");
    return 0;
}

extern "C" int vu_error_labels(const float* p, const float* gt, size_t n, float thr, uint32_t* labels, vu_stream_t st) {
    if (!p || !gt || !labels) return set_error(VU_ERR_INVALID, "error_labels: null");
    error_labels_kernel<<<sm_grid(n), 256, 0, (cudaStream_t)st>>>(p, gt, n, thr, labels);
    VU_CHECK_LAUNCH("error_labels");
    return 0;
}

// From pairs sorted by score descending: out[0] = un-normalised ROC area, out[1] = un-normalised average precision,
// out[2] = number of positives, out[3] = number of negatives. AUROC = out[0] / (out[2] out[3]); AP = out[1] / out[2].
extern "C" int vu_binary_curve_auc(const uint32_t* keys_sorted, const uint32_t* labels_sorted, size_t n, void* workspace,
                                   size_t workspace_bytes, double* out, vu_stream_t st) {
    if (!keys_sorted || !labels_sorted || !workspace || !out || n < 1) return set_error(VU_ERR_INVALID, "binary_curve_auc: bad args");
    if (workspace_bytes < vu_sort_workspace_bytes(n)) return set_error(VU_ERR_INVALID, "binary_curve_auc: workspace too small");
    cudaStream_t s = (cudaStream_t)st;
    const int nb = (int)((n + kSortTile - 1) / kSortTile);
    // [nb + 1] block arrays (the extra slot receives the totals from the exclusive scans) in the histogram area; curve points
    // after the two block-array areas
    uint8_t* hist_area = static_cast<uint8_t*>(workspace) + 4 * align256(n * 4);
    uint32_t* pos_x = reinterpret_cast<uint32_t*>(hist_area);
    uint32_t* end_x = pos_x + (nb + 1);
    uint2* pts = reinterpret_cast<uint2*>(hist_area + align256((size_t)kRadix * nb * 4) + 2 * align256((size_t)nb * 4));
    cudaMemsetAsync(out, 0, 4 * sizeof(double), s);
    cudaMemsetAsync(pos_x, 0, 2 * (size_t)(nb + 1) * 4, s);
    curve_block_counts_kernel<<<nb, kSortThreads, 0, s>>>(keys_sorted, labels_sorted, n, pos_x, end_x);
    scan_u32_single_block_kernel<<<1, 1024, 0, s>>>(pos_x, (size_t)nb + 1);
    scan_u32_single_block_kernel<<<1, 1024, 0, s>>>(end_x, (size_t)nb + 1);
    curve_points_kernel<<<nb, kSortThreads, 0, s>>>(keys_sorted, labels_sorted, n, pos_x, end_x, pts);
    curve_integrate_kernel<<<sm_grid(n), 256, 0, s>>>(pts, end_x + nb, pos_x + nb, n, out);
    VU_CHECK_LAUNCH("binary_curve_auc");
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// Allocation-free variant of vu_sort_pairs_desc_f32 for callers that cannot call cudaMalloc (CUDA graph capture, serving
// loops, harnesses that forbid allocation): the radix sort's workspace (ping-pong key / payload buffers + digit histograms)
// lives in a file-scope device array instead of a caller-provided buffer. Limit: n <= VU_SORT_STATIC_MAX_N = 262 144 pairs (one
// 512 x 512 map); the array is 6.1 MB and is held by every GPU that loads the library until the process exits.
// ALL calls share the array: two calls that overlap on the GPU (different streams, or one device from two threads) race.
// Nothing in it has to start at zero - every word the sort reads was written earlier in the same call (keys by
// make_keys_kernel, every histogram slot by radix_hist_kernel) - so stale contents of the previous call are harmless.
namespace vu {
constexpr size_t kSortStaticMaxN = 262144;
constexpr size_t kSortStaticNb = (kSortStaticMaxN + kSortTile - 1) / kSortTile;
constexpr size_t a256(size_t v) { return (v + 255) / 256 * 256; }
constexpr size_t kSortStaticBytes = 4 * a256(kSortStaticMaxN * 4) + a256((size_t)kRadix * kSortStaticNb * 4) +
                                    2 * a256(kSortStaticNb * 4) + a256(kSortStaticMaxN * 8) + 1024;
static __device__ __align__(256) unsigned char g_sort_scratch[kSortStaticBytes];
}  // namespace vu

extern "C" int vu_sort_pairs_desc_f32_static(const float* score, const uint32_t* vals, size_t n, uint32_t* keys_sorted,
                                             uint32_t* vals_sorted, vu_stream_t st) {
    if (n > kSortStaticMaxN)
        return set_error(VU_ERR_UNSUPPORTED, "sort_pairs_static: n = %zu exceeds the static scratch limit of %zu pairs (use "
                         "vu_sort_pairs_desc_f32 with a caller workspace)", n, (size_t)kSortStaticMaxN);
    // the symbol's name is a host-side stand-in; its device address is looked up once per device, with that device current
    constexpr int kMaxDev = 64;
    static void* dev_ptr[kMaxDev] = {nullptr};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDev)
        return set_error(VU_ERR_CUDA, "sort_pairs_static: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev_ptr[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_sort_scratch);
        if (e != cudaSuccess || p == nullptr)
            return set_error(VU_ERR_CUDA, "sort_pairs_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        dev_ptr[dev] = p;
    }
    if (vu_sort_workspace_bytes(n ? n : 1) > kSortStaticBytes) return set_error(VU_ERR_UNSUPPORTED, "sort_pairs_static: scratch too small");
    return vu_sort_pairs_desc_f32(score, vals, n, keys_sorted, vals_sorted, dev_ptr[dev], kSortStaticBytes, st);
}
