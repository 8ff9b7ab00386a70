// Cost of tcgen05.commit frequency: MMAs (M=128,N=128,K=16, SS) back to back with a commit every G instructions.
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
__global__ void __launch_bounds__(128, 1) k(long long *out, int reps, int G, int wait_mode, int fresh_acc) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bar, ring[8];
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (128 * 32 * 4 + 128 * 32 * 4) / 4; i += 128) ((uint32_t *)smem)[i] = 0;
    if (tid == 0) { ptx::mbar_init(&bar, 1); for (int i = 0; i < 8; ++i) ptx::mbar_init(&ring[i], 1); ptx::fence_mbar_init(); }
    if (warp == 0) ptx::tmem_alloc(&s_tmem, 512);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tb = s_tmem;
    if (tid == 0) {
        const uint32_t a0 = ptx::smem_u32(smem), b0 = a0 + 128 * 32 * 4;
        constexpr uint32_t ID = make_idesc(128, 128);
        uint32_t ncommit = 0;
        long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            const int ks = r & 3;
            // fresh_acc: accumulate flag 0 at the start of every group of 4 (as the head kernel does per K tile of a new
            // output block) vs always accumulating
            ptx::mma_f16_ss(tb + ((r >> 5) & 3) * 128, make_sdesc(a0 + ks * 2 * 128 * 16, 128 * 16),
                            make_sdesc(b0 + ks * 2 * 128 * 16, 128 * 16), ID, fresh_acc ? (r & 31) > 0 : 1);
            if (G > 0 && (r % G) == G - 1) {
                ptx::mma_commit(&ring[ncommit & 7]);
                if (wait_mode == 1 && ncommit >= 4) ptx::mbar_wait(&ring[(ncommit - 4) & 7], ((ncommit - 4) >> 3) & 1);
                ++ncommit;
            }
        }
        ptx::mma_commit(&bar);
        ptx::mbar_wait(&bar, 0);
        long long t1 = clock64();
        out[0] = t1 - t0;
    }
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc(tb, 512);
}
int main() {
    long long *o; cudaMallocManaged(&o, 64);
    size_t smem = 128 * 32 * 8;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int reps = 16384;
    for (int fresh : {0, 1})
        for (int wm : {0, 1})
            for (int G : {0, 32, 8, 4, 2, 1}) {
                k<<<1, 128, smem>>>(o, reps, G, wm, fresh); cudaDeviceSynchronize();
                k<<<1, 128, smem>>>(o, reps, G, wm, fresh); cudaDeviceSynchronize();
                printf("fresh_acc=%d wait_lagging=%d commit every %2d MMAs: %.1f cycles per MMA [%s]\n", fresh, wm, G,
                       (double)o[0] / reps, cudaGetErrorString(cudaGetLastError()));
            }
}
