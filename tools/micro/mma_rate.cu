// How fast does tcgen05.mma (kind::f16, cta_group::1, SS mode, SWIZZLE_NONE K-major operands) run back to back?
// Shapes: M=128, N in {128, 256}; K=16 per instruction; operands as laid out by mlp_tc.cu (LBO = rows*16, SBO = 128).
#include <cstdio>
#include <cuda_runtime.h>
#include "../../deepclr_b200/csrc/ptx.cuh"
using namespace dclr;
__host__ __device__ constexpr uint32_t make_idesc(int M, int N) {
    return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ uint64_t make_sdesc(uint32_t smem_addr, uint32_t lbo_bytes) {
    return (uint64_t)((smem_addr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
           ((uint64_t)(128u >> 4) << 32) | (1ull << 46);
}
template <int N, int KSTEPS>
__global__ void __launch_bounds__(128, 1) k(long long *out, int reps) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (128 * 16 * 2 * KSTEPS + N * 16 * 2 * KSTEPS) / 4; i += 128) ((uint32_t *)smem)[i] = 0;
    if (tid == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
    if (warp == 0) ptx::tmem_alloc(&s_tmem, 512);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tb = s_tmem;
    if (tid == 0) {
        const uint32_t a0 = ptx::smem_u32(smem), b0 = a0 + 128 * 32 * KSTEPS;
        constexpr uint32_t ID = make_idesc(128, N);
        long long t0 = clock64();
        for (int r = 0; r < reps; ++r)
#pragma unroll
            for (int ks = 0; ks < KSTEPS; ++ks)
                ptx::mma_f16_ss(tb + (r & 1) * N % 512, make_sdesc(a0 + ks * 2 * 128 * 16, 128 * 16),
                                make_sdesc(b0 + ks * 2 * N * 16, N * 16), ID, ks > 0);
        ptx::mma_commit(&bar);
        ptx::mbar_wait(&bar, 0);
        long long t1 = clock64();
        out[blockIdx.x] = t1 - t0;
    }
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc(tb, 512);
}
template <int N, int KSTEPS>
void run(const char *name, int grid) {
    long long *o; cudaMallocManaged(&o, 148 * 8);
    size_t smem = 128 * 32 * KSTEPS + N * 32 * KSTEPS;
    cudaFuncSetAttribute(k<N, KSTEPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int reps = 2000;
    k<N, KSTEPS><<<grid, 128, smem>>>(o, reps); cudaDeviceSynchronize();
    k<N, KSTEPS><<<grid, 128, smem>>>(o, reps); cudaDeviceSynchronize();
    double cyc = (double)o[0] / (reps * KSTEPS);
    printf("%s grid=%d: %.1f cycles per MMA (M=128,N=%d,K=16) -> %.0f MAC/cycle/SM  [%s]\n", name, grid, cyc, N,
           128.0 * N * 16 / cyc, cudaGetErrorString(cudaGetLastError()));
    cudaFree(o);
}
int main() {
    run<128, 4>("N128", 1); run<256, 4>("N256", 1); run<128, 4>("N128", 148); run<256, 4>("N256", 148);
    run<64, 4>("N64", 1);
}
