// Internal helpers shared by the translation units of libvaeunet_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/vaeunet_b200.h"

namespace vu {

int set_error(int code, const char* fmt, ...);
int num_sms();

#define VU_CHECK_LAUNCH(name)                                                              \
    do {                                                                                   \
        cudaError_t e__ = cudaGetLastError();                                              \
        if (e__ != cudaSuccess)                                                            \
            return ::vu::set_error(VU_ERR_CUDA, "%s launch: %s", name, cudaGetErrorString(e__)); \
    } while (0)

}  // namespace vu
