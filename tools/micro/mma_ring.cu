// Round-trip of the producer/consumer ring the head kernel uses, without any data movement:
//   MMA thread : wait full[st] -> 4 x tcgen05.mma (128x128x16) -> tcgen05.commit -> empty[st]
//   producer   : wait empty[st] -> arrive full[st]
// cycles per tile as a function of the ring depth tells the commit -> mbarrier -> waiter latency.
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
template <int STAGES, int MMAS>
__global__ void __launch_bounds__(96, 1) k(long long *out, int tiles, int mode) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bar, full[STAGES], empty[STAGES];
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < 32768 / 4; i += 96) ((uint32_t *)smem)[i] = 0;
    if (tid == 0) {
        ptx::mbar_init(&bar, 1);
        for (int i = 0; i < STAGES; ++i) { ptx::mbar_init(&full[i], 1); ptx::mbar_init(&empty[i], 1); }
        ptx::fence_mbar_init();
    }
    if (warp == 0) ptx::tmem_alloc(&s_tmem, 512);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tb = s_tmem;
    if (warp == 0 && (mode & 256)) {
        // converged warp, bookkeeping tucked under the tile's own MMAs: after the first J MMAs of tile t, commit tile
        // t-1's stage (the commit also covers those J MMAs: harmless, the stage is released ~J x 64 cycles later) and
        // wait for tile t+1's stage; then the remaining MMAs.  J = mode >> 9.
        const int J = mode >> 9;
        const uint32_t a0 = ptx::smem_u32(smem), b0 = a0 + 16384;
        constexpr uint32_t ID = make_idesc(128, 128);
        uint32_t st = 0, ph = 0;
        auto mma = [&](uint32_t d, int ks) {
            const uint64_t ad = make_sdesc(a0 + (ks & 3) * 4096, 2048), bd = make_sdesc(b0 + (ks & 3) * 4096, 2048);
            asm volatile(
                "{\n\t.reg .pred p, e;\n\t"
                "elect.sync _|e, 0xffffffff;\n\t"
                "setp.ne.b32 p, %4, 0;\n\t"
                "@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd), "r"(ID), "r"(1u) : "memory");
        };
        long long t0 = clock64();
        ptx::mbar_wait(&full[0], 0);
        ptx::tc_fence_after();
        for (int t = 0; t < tiles; ++t) {
            const uint32_t d = tb + (t & 3) * 128;
            uint32_t nst = st + 1, nph = ph;
            if (nst == STAGES) { nst = 0; nph ^= 1; }
#pragma unroll
            for (int ks = 0; ks < MMAS; ++ks) {
                if (ks == J) {
                    if (t > 0) {
                        const uint32_t pst = st == 0 ? STAGES - 1 : st - 1;
                        asm volatile(
                            "{\n\t.reg .pred e;\n\t"
                            "elect.sync _|e, 0xffffffff;\n\t"
                            "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(ptx::smem_u32(&empty[pst])) : "memory");
                    }
                    if (t + 1 < tiles) { ptx::mbar_wait(&full[nst], nph); ptx::tc_fence_after(); }
                }
                mma(d, ks);
            }
            st = nst; ph = nph;
        }
        if (lane == 0) {
            const uint32_t pst = st == 0 ? STAGES - 1 : st - 1;
            ptx::mma_commit(&empty[pst]);
            ptx::mma_commit(&bar); ptx::mbar_wait(&bar, 0);
        }
        __syncwarp();
        long long t1 = clock64();
        if (lane == 0) out[0] = t1 - t0;
    } else if (warp == 0 && (mode & 128)) {
        // converged warp: every lane runs the loop, one elected lane issues
        const uint32_t a0 = ptx::smem_u32(smem), b0 = a0 + 16384;
        constexpr uint32_t ID = make_idesc(128, 128);
        uint32_t st = 0, ph = 0;
        long long t0 = clock64();
        for (int t = 0; t < tiles; ++t) {
            if (!(mode & 1)) ptx::mbar_wait(&full[st], ph);
            if (!(mode & 2)) ptx::tc_fence_after();
            const uint32_t d = tb + (t & 3) * 128;
#pragma unroll
            for (int ks = 0; ks < MMAS; ++ks) {
                const uint64_t ad = make_sdesc(a0 + (ks & 3) * 4096, 2048), bd = make_sdesc(b0 + (ks & 3) * 4096, 2048);
                asm volatile(
                    "{\n\t.reg .pred p, e;\n\t"
                    "elect.sync _|e, 0xffffffff;\n\t"
                    "setp.ne.b32 p, %4, 0;\n\t"
                    "@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd), "r"(ID), "r"(1u) : "memory");
            }
            if (!(mode & 32)) {
                asm volatile(
                    "{\n\t.reg .pred e;\n\t"
                    "elect.sync _|e, 0xffffffff;\n\t"
                    "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(ptx::smem_u32(&empty[st])) : "memory");
            }
            if (++st == STAGES) { st = 0; ph ^= 1; }
        }
        if (lane == 0) { ptx::mma_commit(&bar); ptx::mbar_wait(&bar, 0); }
        __syncwarp();
        long long t1 = clock64();
        if (lane == 0) out[0] = t1 - t0;
    } else if (warp == 0 && lane == 0) {
        const uint32_t a0 = ptx::smem_u32(smem), b0 = a0 + 16384;
        constexpr uint32_t ID = make_idesc(128, 128);
        uint64_t ad[4], bd[4];
        for (int ks = 0; ks < 4; ++ks) { ad[ks] = make_sdesc(a0 + ks * 4096, 2048); bd[ks] = make_sdesc(b0 + ks * 4096, 2048); }
        uint32_t st = 0, ph = 0;
        long long t0 = clock64();
        for (int t = 0; t < tiles; ++t) {
            // mode bit0: skip the full wait; bit1: skip the fence; bit2: commit only every 4th tile (to a dummy barrier)
            if (!(mode & 1)) ptx::mbar_wait(&full[st], ph);
            if (!(mode & 2)) ptx::tc_fence_after();
#pragma unroll
            for (int ks = 0; ks < MMAS; ++ks) ptx::mma_f16_ss(tb + ((mode & 16) ? (t & 1) : (t & 3)) * 128, ad[ks & 3], bd[ks & 3], ID, (mode & 8) ? (uint32_t)(ks > 0) : 1u);
            if (mode & 32) { /* no commit in the loop */ }
            else if (mode & 64) { if ((t & 63) == 63) ptx::mma_commit(&empty[st]); }
            else if (!(mode & 4)) ptx::mma_commit(&empty[st]);
            else if ((t & 3) == 3) ptx::mma_commit(&empty[st]);
            if (++st == STAGES) { st = 0; ph ^= 1; }
        }
        ptx::mma_commit(&bar);
        ptx::mbar_wait(&bar, 0);
        long long t1 = clock64();
        out[0] = t1 - t0;
    } else if (warp == 1 && lane == 0) {
        uint32_t st = 0, ph = 0;
        for (int t = 0; t < tiles && !(mode & 1); ++t) {
            ptx::mbar_wait(&empty[st], ph ^ 1);
            ptx::mbar_arrive(&full[st]);
            if (++st == STAGES) { st = 0; ph ^= 1; }
        }
    }
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc(tb, 512);
}
template <int STAGES, int MMAS>
void run(int mode = 0) {
    long long *o; cudaMallocManaged(&o, 64);
    cudaFuncSetAttribute(k<STAGES, MMAS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
    const int tiles = 4000;
    for (int i = 0; i < 2; ++i) { k<STAGES, MMAS><<<1, 96, 32768>>>(o, tiles, mode); cudaDeviceSynchronize(); }
    printf("mode %d ring depth %2d, %d MMAs per tile: %.0f cycles per tile (MMA time %d) [%s]\n", mode, STAGES, MMAS, (double)o[0] / tiles,
           MMAS * 64, cudaGetErrorString(cudaGetLastError()));
    fflush(stdout);
    cudaFree(o);
}
int main() {
    run<5, 4>(0);            // single-lane issue (as the kernels do today), full ring
    run<5, 4>(128);          // converged warp + elect, full ring
    run<5, 4>(128 | 1 | 2);  // converged warp, no wait / fence
    run<5, 4>(128 | 1 | 2 | 32);
    run<2, 4>(128);
    run<5, 8>(128);
    // which part of the per-tile bookkeeping is exposed (7 MMAs per tile = the head kernel's 112-wide K tiles)
    run<3, 7>(128);                  // wait + fence + commit
    run<3, 7>(128 | 2);              // wait + commit
    run<3, 7>(128 | 1);              // fence + commit
    run<3, 7>(128 | 1 | 2);          // commit only
    run<3, 7>(128 | 1 | 32);         // fence only
    run<3, 7>(128 | 1 | 2 | 32);     // nothing
    for (int J = 0; J <= 4; ++J) run<3, 7>(256 | (J << 9));     // bookkeeping after the J-th MMA of the tile
    for (int J = 1; J <= 3; ++J) run<3, 4>(256 | (J << 9));
    run<5, 7>(128);
    run<8, 7>(128);
    run<3, 14>(128);
    run<3, 28>(128);
}
