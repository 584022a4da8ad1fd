// Error plumbing + device checks of the C ABI.
#include "vu_internal.h"
#include <cstdarg>
#include <cstdio>

namespace vu {

static thread_local char g_err[512] = "";

int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int num_sms() {
    static int sms[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (sms[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        sms[dev] = v;
    }
    return sms[dev];
}

}  // namespace vu

extern "C" const char* vu_last_error(void) { return vu::g_err; }
extern "C" int vu_version(void) { return 100; }

extern "C" int vu_device_check(void) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return vu::set_error(VU_ERR_NO_DEVICE, "no CUDA device: %s", cudaGetErrorString(e));
    int major = 0, minor = 0;
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
    if (major != 10)
        return vu::set_error(VU_ERR_NO_DEVICE, "device is sm_%d%d; this library is sm_100a only (no fallback)", major, minor);
    return 0;
}
