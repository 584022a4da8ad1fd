// Pieces shared by the tcgen05 convolution kernels: 16-bit pack/unpack for the epilogues and the
// host-side TMA descriptor (CUtensorMap) encoder.
#pragma once
#include "vu_ptx.cuh"
#include "vu_internal.h"
#include "vu_half.cuh"
#include <cuda_bf16.h>
#include <cuda_fp16.h>

namespace vu {

constexpr int kMaxTapsCommon = 9;

__device__ __forceinline__ uint32_t pack2(float a, float b, int dtype) {
    if (dtype == 0) {
        __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&v);
    } else {
        return cvt_f16x2_sat(a, b);      // saturating (vu_half.cuh): no silent inf in fp16 activations
    }
}
__device__ __forceinline__ float2 unpack2(uint32_t u, int dtype) {
    if (dtype == 0) {
        __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
        return __bfloat1622float2(v);
    } else {
        __half2 v = *reinterpret_cast<__half2*>(&u);
        return __half22float2(v);
    }
}

// Column sums of a 32 x 32 tile held one row per lane (a[i] = this lane's value for column i): a butterfly that halves the
// number of live columns per lane at every step (16+8+4+2+1 = 31 shuffles). Returns, in lane l, the sum over the 32 lanes
// of column l. Used by the conv epilogues to accumulate per-channel BatchNorm statistics of the tile they just produced.
__device__ __forceinline__ float warp_column_sums32(float (&a)[32], int lane) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        const bool hi = (lane & o) != 0;
#pragma unroll
        for (int i = 0; i < o; ++i) {
            const float send = hi ? a[i] : a[i + o];
            const float keep = hi ? a[i + o] : a[i];
            a[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    return a[0];
}

// ------------------------------------------------------------------ BatchNorm statistics in the conv epilogues
// Per-channel sum / sum of squares of a 32-pixel x 32-channel chunk held one ROW per lane as 16 packed 16-bit pairs (the
// values exactly as stored). The chunk is transposed through a 2 KB per-warp scratch (rows of 64 B, 16-byte pieces XOR-
// swizzled by (row >> 1) & 3: conflict-free both ways): lane = (h, cp) then sums the column pair (2cp, 2cp+1) over the 16 rows of
// parity h, one shuffle combines the two halves, and lanes 0..15 add the result into this warp's OWN slice of the per-CTA
// partial arrays with plain read-modify-writes - each (TMEM lane quarter, 32-column chunk) belongs to exactly one warp, so
// there are no atomics and the summation order is fixed. ~120 instructions per chunk; the first version (two 31-shuffle
// butterflies + fp64 shared-memory atomics, which compile to CAS loops) cost ~330 and made the training convs epilogue-bound.
constexpr int kEpiStatMaxCout = 512;
constexpr int kEpiStatBytes = 4 * 2 * kEpiStatMaxCout * 4;     // [4 lane quarters][sum, sumsq][Cout] fp32
constexpr int kEpiScratchBytes = 32 * 64;                       // per epilogue warp
__device__ __forceinline__ void epi_stats_chunk(const uint32_t (&outw)[16], bool valid, int lane, int dtype, uint8_t* scratch,
                                                float* part_sum, float* part_sq) {
    uint4* srow = reinterpret_cast<uint4*>(scratch + lane * 64);
#pragma unroll
    for (int j = 0; j < 4; ++j)
        srow[j ^ ((lane >> 1) & 3)] = valid ? make_uint4(outw[4 * j], outw[4 * j + 1], outw[4 * j + 2], outw[4 * j + 3])
                                     : make_uint4(0, 0, 0, 0);
    __syncwarp();
    const int h = lane >> 4, cp = lane & 15;
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int r = 2 * k + h;
        const uint32_t w = *reinterpret_cast<const uint32_t*>(scratch + r * 64 + (((cp >> 2) ^ (k & 3)) << 4) + ((cp & 3) << 2));
        const float2 f = unpack2(w, dtype);
        s0 += f.x; s1 += f.y;
        q0 = fmaf(f.x, f.x, q0); q1 = fmaf(f.y, f.y, q1);
    }
    s0 += __shfl_xor_sync(0xffffffffu, s0, 16); s1 += __shfl_xor_sync(0xffffffffu, s1, 16);
    q0 += __shfl_xor_sync(0xffffffffu, q0, 16); q1 += __shfl_xor_sync(0xffffffffu, q1, 16);
    __syncwarp();                                   // the scratch is rewritten by this warp's next chunk
    if (lane < 16) {
        float2* ps = reinterpret_cast<float2*>(part_sum + 2 * cp);
        float2* pq = reinterpret_cast<float2*>(part_sq + 2 * cp);
        float2 a = *ps, b = *pq;
        a.x += s0; a.y += s1; b.x += q0; b.y += q1;
        *ps = a; *pq = b;
    }
}
// After the block-wide barrier at the end of a kernel: combine the `nparts` private slices in a fixed order in fp64 and flush
// this CTA's partials with one fp64 atomic per channel and statistic.
__device__ __forceinline__ void epi_stats_flush(const float* s_stat, double* stat_sums, int cout, int tid, int nthreads,
                                                int stride = kEpiStatMaxCout, int nparts = 4) {
    for (int c = tid; c < cout; c += nthreads) {
        double a = 0.0, b = 0.0;
        for (int q = 0; q < nparts; ++q) {
            a += (double)s_stat[(q * 2) * stride + c];
            b += (double)s_stat[(q * 2 + 1) * stride + c];
        }
        if (a != 0.0 || b != 0.0) {
            atomicAdd(&stat_sums[c], a);
            atomicAdd(&stat_sums[cout + c], b);
        }
    }
}

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);

static inline PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* f = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(f);
    }
    return fn;
}

static inline int encode_map(CUtensorMap* m, int dtype, int rank, const void* base, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box) {
    PFN_encodeTiled fn = get_encode_fn();
    if (!fn) return set_error(VU_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    CUresult r = fn(m, dtype == 0 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
                    rank, const_cast<void*>(base), dims, strides_bytes, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(VU_ERR_CUDA, "cuTensorMapEncodeTiled failed (CUresult %d)", (int)r);
    return 0;
}


// vu_conv_halo.cu: 3x3 s1 p1 convs with Cout in {64,128} on wide maps (smem tap reuse). Returns 1 when it
// took the launch, 0 when the shape is not its case, < 0 on error.
int conv3x3_halo_launch(const vu_conv_args* a, cudaStream_t stream);
// vu_conv_flat.cu: 3x3 s1 p1 convs on narrow maps (W <= 64), tap reuse over a flat padded pixel index. Same return values.
int conv3x3_flat_launch(const vu_conv_args* a, cudaStream_t stream);

}  // namespace vu
