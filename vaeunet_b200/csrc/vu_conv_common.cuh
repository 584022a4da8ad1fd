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

}  // namespace vu
