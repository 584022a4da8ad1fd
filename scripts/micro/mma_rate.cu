// Micro-benchmark: issue rate of tcgen05.mma kind::f16 (M=128, K=16) from smem operands, per N and per
// number of accumulators used round-robin. One CTA per SM, no TMA (operands are whatever smem holds).
#include "../../vaeunet_b200/csrc/vu_ptx.cuh"
#include <cstdio>
using namespace vu;

template <int N, int NACC, int AOFF>
__global__ void __launch_bounds__(128, 1) k_rate(long long* out, int iters) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_barrier_init(); }
    if (warp == 0) { tmem_alloc<512>(smem_u32(&slot)); tmem_relinquish(); }
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tmem = slot;
    if (warp == 0 && lane == 0) {
        const uint32_t idesc = umma_idesc_f16(128, N, 0);
        const uint32_t a0 = base + AOFF * 128, b0 = base + 96 * 1024;
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int j = 0; j < NACC; ++j)
                    umma_f16_ss(tmem + j * N, umma_desc_k_sw128(a0 + j * 16640) + 2u * k, umma_desc_k_sw128(b0) + 2u * k, idesc, 1);
        }
        umma_commit(smem_u32(&bar));
        mbar_wait(smem_u32(&bar), 0);
        long long t1 = clock64();
        if (blockIdx.x == 0) out[0] = t1 - t0;
    }
    tc_fence_before(); __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

template <int N, int NACC, int AOFF> void run(const char* name) {
    long long* d; cudaMalloc(&d, 8);
    const int smem = 200 * 1024, iters = 2000;
    cudaFuncSetAttribute(k_rate<N, NACC, AOFF>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int grid : {1, 148}) {
        k_rate<N, NACC, AOFF><<<grid, 128, smem>>>(d, iters);
        cudaError_t e = cudaDeviceSynchronize();
        long long h = 0; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        printf("%-28s grid %3d: %7.1f clk / MMA (%s)\n", name, grid, (double)h / (iters * 4.0 * NACC), cudaGetErrorString(e));
    }
    cudaFree(d);
}

int main() {
    run<64, 1, 0>("N=64  1 acc aligned");
    run<64, 2, 0>("N=64  2 acc aligned");
    run<64, 4, 0>("N=64  4 acc aligned");
    run<64, 4, 3>("N=64  4 acc start+3rows");
    run<128, 1, 0>("N=128 1 acc aligned");
    run<128, 2, 0>("N=128 2 acc aligned");
    run<128, 2, 3>("N=128 2 acc start+3rows");
    run<256, 1, 0>("N=256 1 acc aligned");
    run<256, 2, 0>("N=256 2 acc aligned");
    return 0;
}
