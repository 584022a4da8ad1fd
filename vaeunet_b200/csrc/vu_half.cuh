// 16-bit storage helpers: dtype 0 = bf16, 1 = fp16 (runtime switch; all math is fp32).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace vu {

// fp16 conversions SATURATE to +-65504 (cvt.rn.satfinite: one F2FP instruction, same cost as the plain conversion): an
// activation beyond the fp16 range of a real checkpoint becomes the largest finite value instead of a silent inf that would
// turn every downstream BatchNorm / sigmoid into NaN. NaN stays NaN. bf16 has the fp32 exponent range and needs none.
__device__ __forceinline__ uint32_t cvt_f16x2_sat(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
__device__ __forceinline__ uint16_t f2h(float v, int dtype) {
    if (dtype == 0) { __nv_bfloat16 b = __float2bfloat16_rn(v); return *reinterpret_cast<uint16_t*>(&b); }
    return static_cast<uint16_t>(cvt_f16x2_sat(v, 0.f) & 0xffffu);
}
__device__ __forceinline__ float h2f(uint16_t u, int dtype) {
    if (dtype == 0) return __uint_as_float(static_cast<uint32_t>(u) << 16);
    __half h = *reinterpret_cast<__half*>(&u); return __half2float(h);
}
__device__ __forceinline__ uint32_t pack2h(float a, float b, int dtype) {
    if (dtype == 0) { __nv_bfloat162 v = __floats2bfloat162_rn(a, b); return *reinterpret_cast<uint32_t*>(&v); }
    return cvt_f16x2_sat(a, b);
}
__device__ __forceinline__ void unpack2h(uint32_t u, float& a, float& b, int dtype) {
    if (dtype == 0) { a = __uint_as_float(u << 16); b = __uint_as_float(u & 0xffff0000u); return; }
    __half2 v = *reinterpret_cast<__half2*>(&u); float2 f = __half22float2(v); a = f.x; b = f.y;
}
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8], int dtype) {
    unpack2h(v.x, f[0], f[1], dtype); unpack2h(v.y, f[2], f[3], dtype);
    unpack2h(v.z, f[4], f[5], dtype); unpack2h(v.w, f[6], f[7], dtype);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8], int dtype) {
    return make_uint4(pack2h(f[0], f[1], dtype), pack2h(f[2], f[3], dtype), pack2h(f[4], f[5], dtype),
                      pack2h(f[6], f[7], dtype));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace vu
