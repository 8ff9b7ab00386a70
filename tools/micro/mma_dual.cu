// Two MMA-issuing warps on one SM, each with its own producer/consumer ring (wait full -> MMAS x tcgen05.mma -> commit ->
// empty) and its own TMEM accumulator columns: does the bookkeeping of one warp hide under the MMAs of the other?
// (One warp alone: 586 cycles per 7-MMA tile against 448 of tensor time, wherever the wait / commit sit: mma_ring.cu.)
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
template <int STAGES, int MMAS, int N>
__global__ void __launch_bounds__(128, 1) k(long long *out, int tiles, int issuers, int acc_mask, int prod_warp) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bar[2], full[2][STAGES], empty[2][STAGES];
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < 49152 / 4; i += blockDim.x) ((uint32_t *)smem)[i] = 0;
    if (tid == 0) {
        for (int w = 0; w < 2; ++w) {
            ptx::mbar_init(&bar[w], 1);
            for (int i = 0; i < STAGES; ++i) { ptx::mbar_init(&full[w][i], 1); ptx::mbar_init(&empty[w][i], 1); }
        }
        ptx::fence_mbar_init();
    }
    if (warp == 0) ptx::tmem_alloc(&s_tmem, 512);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tb = s_tmem;
    if (warp < issuers) {
        const uint32_t a0 = ptx::smem_u32(smem), b0 = a0 + 16384;
        constexpr uint32_t ID = make_idesc(128, N);
        uint32_t st = 0, ph = 0;
        long long t0 = clock64();
        for (int t = 0; t < tiles; ++t) {
            ptx::mbar_wait(&full[warp][st], ph);
            ptx::tc_fence_after();
            const uint32_t d = issuers == 1 ? tb + (N == 128 ? (t & acc_mask) * 128 : (t & acc_mask & 1) * 256) : tb + warp * 256 + (N == 128 ? (t & 1) * 128 : 0);
#pragma unroll
            for (int ks = 0; ks < MMAS; ++ks) {
                const uint64_t ad = make_sdesc(a0 + (ks & 3) * 4096, 2048), bd = make_sdesc(b0 + (ks & 3) * 2 * N * 16, N * 16);
                asm volatile(
                    "{\n\t.reg .pred p, e;\n\t"
                    "elect.sync _|e, 0xffffffff;\n\t"
                    "setp.ne.b32 p, %4, 0;\n\t"
                    "@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd), "r"(ID), "r"(1u) : "memory");
            }
            asm volatile(
                "{\n\t.reg .pred e;\n\t"
                "elect.sync _|e, 0xffffffff;\n\t"
                "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(ptx::smem_u32(&empty[warp][st])) : "memory");
            if (++st == STAGES) { st = 0; ph ^= 1; }
        }
        if (lane == 0) { ptx::mma_commit(&bar[warp]); ptx::mbar_wait(&bar[warp], 0); }
        __syncwarp();
        long long t1 = clock64();
        if (lane == 0) out[warp] = t1 - t0;
    } else if (((issuers == 2 && warp >= 2) || (issuers == 1 && warp == prod_warp)) && lane == 0) {
        const int w = issuers == 2 ? warp - 2 : 0;
        uint32_t st = 0, ph = 0;
        for (int t = 0; t < tiles; ++t) {
            ptx::mbar_wait(&empty[w][st], ph ^ 1);
            ptx::mbar_arrive(&full[w][st]);
            if (++st == STAGES) { st = 0; ph ^= 1; }
        }
    }
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc(tb, 512);
}
template <int STAGES, int MMAS, int N>
void run(int issuers, int acc_mask = 1, int prod_warp = 2, int threads = 128) {
    long long *o; cudaMallocManaged(&o, 64);
    cudaFuncSetAttribute(k<STAGES, MMAS, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152);
    const int tiles = 4000;
    for (int i = 0; i < 2; ++i) { o[0] = o[1] = 0; k<STAGES, MMAS, N><<<1, threads, 49152>>>(o, tiles, issuers, acc_mask, prod_warp); cudaDeviceSynchronize(); }
    const double cyc = (double)(o[0] > o[1] ? o[0] : o[1]) / (tiles * issuers);
    printf("acc_mask %d prod_warp %d threads %d | issuers %d, ring depth %d, %d MMAs (N=%d) per tile: %.0f cycles per tile (tensor time %d) [%s]\n", acc_mask, prod_warp, threads, issuers, STAGES, MMAS, N,
           cyc, MMAS * N / 2, cudaGetErrorString(cudaGetLastError()));
    cudaFree(o);
}
int main() {
    run<3, 7, 128>(1, 3); run<3, 7, 128>(1, 1, 1); run<3, 7, 128>(1, 3, 1); run<3, 7, 128>(1, 3, 1, 96); run<3, 7, 128>(1, 1, 1, 96); run<3, 7, 128>(1, 0);
    run<3, 7, 128>(1); run<3, 7, 128>(2);
    run<2, 5, 128>(1); run<2, 5, 128>(2);
    run<2, 4, 128>(1); run<2, 4, 128>(2);
    run<3, 7, 256>(1); run<3, 7, 256>(2);
    run<3, 3, 256>(1); run<3, 3, 256>(2);
}
