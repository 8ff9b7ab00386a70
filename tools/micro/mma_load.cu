// Does a concurrent stream of TMA bulk copies into shared memory slow tcgen05.mma (SS mode) down?
// One warp issues MMAs back to back on resident operands, another keeps STAGES x 16 KB bulk copies from L2 in flight
// into other shared memory.  Reports cycles per MMA with and without the copy stream, for N = 128 and N = 256.
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
#ifndef STAGES
#define STAGES 5
#endif
constexpr int TILE = 16384, KSTEPS = 4;
template <int N>
__global__ void __launch_bounds__(96, 1) k(const uint8_t *src, long long *out, int reps, int with_copy, int spread) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bar, full[STAGES];
    __shared__ uint32_t s_tmem;
    __shared__ volatile int stop;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t opbytes = 128 * 32 * KSTEPS + N * 32 * KSTEPS;
    for (int i = tid; i < (int)opbytes / 4; i += 96) ((uint32_t *)smem)[i] = 0;
    if (tid == 0) { ptx::mbar_init(&bar, 1); for (int i = 0; i < STAGES; ++i) ptx::mbar_init(&full[i], 1); stop = 0; ptx::fence_mbar_init(); }
    if (warp == 0) ptx::tmem_alloc(&s_tmem, 512);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tb = s_tmem;
    uint8_t *ring = smem + ((opbytes + 1023) / 1024) * 1024;
    if (warp == 0 && lane == 0) {
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
        out[blockIdx.x * 2] = t1 - t0;
        stop = 1;
    } else if (warp == 1 && lane == 0 && with_copy) {
        long long n = 0;
        for (uint32_t it = 0; !stop; ++it) {
            const int st = it % STAGES;
            if (it >= STAGES) ptx::mbar_wait(&full[st], ((it / STAGES) - 1) & 1);
            ptx::mbar_expect_tx(&full[st], TILE);
            ptx::bulk_g2s(ring + st * TILE, src + (size_t)((it * 7 + (spread ? blockIdx.x : 0)) % 128) * TILE, TILE, &full[st]);
            ++n;
        }
        out[blockIdx.x * 2 + 1] = n;
        // drain
        for (int st = 0; st < STAGES; ++st) { /* outstanding copies complete before the CTA exits via the barrier waits below */ }
    }
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc(tb, 512);
}
template <int N>
void run(const uint8_t *src, int with_copy, int spread = 1, int grid = 148) {
    long long *o; cudaMallocManaged(&o, 148 * 16);
    size_t smem = 128 * 32 * KSTEPS + N * 32 * KSTEPS + 1024 + STAGES * TILE;
    cudaFuncSetAttribute(k<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int reps = 20000;
    for (int i = 0; i < 2; ++i) { o[1] = 0; k<N><<<grid, 96, smem>>>(src, o, reps, with_copy, spread); cudaDeviceSynchronize(); }
    const double cyc = (double)o[0] / (reps * KSTEPS);
    printf("stages=%d grid=%d spread=%d N=%d copy=%d: %.1f cycles per MMA (ideal %d); copies landed per CTA: %lld = %.1f B/cycle  [%s]\n", STAGES, grid, spread, N, with_copy,
           cyc, N / 2, o[1], (double)o[1] * TILE / (double)o[0], cudaGetErrorString(cudaGetLastError()));
    cudaFree(o);
}
int main() {
    uint8_t *src; cudaMalloc(&src, 128 * TILE); cudaMemset(src, 0, 128 * TILE);
    run<128>(src, 0); run<128>(src, 1); run<128>(src, 1, 1, 1); run<128>(src, 1, 1, 16); run<128>(src, 1, 1, 74);
}
