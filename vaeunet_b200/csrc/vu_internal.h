// Internal helpers shared by the translation units of libvaeunet_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/vaeunet_b200.h"

namespace vu {

int set_error(int code, const char* fmt, ...);
int num_sms();
// vu_small_cin_mma.cu: tensor-core first conv (Cin = 3); 0 = launched, -1 = shape not covered
int launch_small_cin_mma(int K, int S, const float* x, const float* w, const float* scale, const float* bias, void* y, int N,
                         int Cin, int H, int W, int Cout, int pad, int y_ctot, int y_coff, int relu, int dtype,
                         cudaStream_t stream);

#define VU_CHECK_LAUNCH(name)                                                              \
    do {                                                                                   \
        cudaError_t e__ = cudaGetLastError();                                              \
        if (e__ != cudaSuccess)                                                            \
            return ::vu::set_error(VU_ERR_CUDA, "%s launch: %s", name, cudaGetErrorString(e__)); \
    } while (0)

}  // namespace vu
