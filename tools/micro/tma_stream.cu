// How fast can one SM pull an L2-resident array into shared memory with cp.async.bulk (the head kernel's weight stream)?
// P producer lanes (one per warp), each with S stages of TILE bytes in flight; source = a 2 MB array (L2 resident), read
// round-robin.  Reports bytes per cycle per SM for grid = 1 and 148.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../deepclr_b200/csrc/ptx.cuh"
using namespace dclr;
__global__ void __launch_bounds__(128, 1) k(const uint8_t *src, long long *out, int copies, int P, int S, int tile, int src_bytes) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t full[4][16];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) { for (int p = 0; p < 4; ++p) for (int i = 0; i < 16; ++i) ptx::mbar_init(&full[p][i], 1); ptx::fence_mbar_init(); }
    __syncthreads();
    if (warp < P && lane == 0) {
        uint8_t *ring = smem + (size_t)warp * S * tile;
        const int ntiles = src_bytes / tile;
        long long t0 = clock64();
        for (int it = 0; it < copies; ++it) {
            const int st = it % S;
            if (it >= S) ptx::mbar_wait(&full[warp][st], ((it / S) - 1) & 1);
            ptx::mbar_expect_tx(&full[warp][st], tile);
            ptx::bulk_g2s(ring + st * tile, src + (size_t)((it * P + warp + blockIdx.x * 7) % ntiles) * tile, tile, &full[warp][st]);
        }
        for (int it = copies; it < copies + S; ++it) {
            const int st = it % S;
            if (it >= S) ptx::mbar_wait(&full[warp][st], ((it / S) - 1) & 1);
        }
        long long t1 = clock64();
        out[blockIdx.x * 4 + warp] = t1 - t0;
    }
}
int main() {
    const int src_bytes = 2 << 20;
    uint8_t *src; cudaMalloc(&src, src_bytes); cudaMemset(src, 0, src_bytes);
    long long *o; cudaMallocManaged(&o, 148 * 4 * 8);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int copies = 4000;
    for (int grid : {1, 148})
        for (int tile : {4096, 16384, 28672, 65536})
            for (int P : {1, 2, 4})
                for (int S : {1, 2, 3, 6}) {
                    if ((size_t)P * S * tile > 192 * 1024) continue;
                    for (int i = 0; i < 2; ++i) { k<<<grid, 128, (size_t)P * S * tile>>>(src, o, copies, P, S, tile, src_bytes); cudaDeviceSynchronize(); }
                    long long mx = 0;
                    for (int b = 0; b < grid; ++b) for (int p = 0; p < P; ++p) if (o[b * 4 + p] > mx) mx = o[b * 4 + p];
                    printf("grid %3d tile %5d B, %d producers x %d stages: %.1f B/cycle/SM [%s]\n", grid, tile, P, S,
                           (double)copies * P * tile / (double)mx, cudaGetErrorString(cudaGetLastError()));
                }
}
