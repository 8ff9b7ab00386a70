// Row gather (128-byte rows of a 16 MB L2-resident table, random row order) global -> shared through the LSU, all SMs:
// which lane mapping / instruction gets closest to the L2 cap (~42 B/cycle/SM)?
//   ./lsu_gather <mode> <warps> <depth> [iters]
// mode 0: thread = row, 8 x LDG.128 + STS.128 (32 lines per instruction)
// mode 1: 8 rows x 4 chunks per instruction, LDGSTS.cg.128   (8 lines, conflict-free stores)
// mode 2: 4 rows x 8 chunks per instruction, LDGSTS.cg.128   (4 full lines)
// mode 3: as 1 with LDG.128 + STS.128
// mode 4: as 1 with LDGSTS.ca
// depth = cp.async groups (or register batches) a warp keeps in flight.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "../../deepclr_b200/csrc/ptx.cuh"
using namespace dclr;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int ROWS_PER_WARP_STEP = 32;          // a warp moves 32 rows (4 KB) per step in every mode

__global__ void __launch_bounds__(1024, 1)
k(const uint4 *tab, const int *rows, int steps, int mode, int depth, long long *cyc, uint32_t *sink) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
    uint8_t *buf = smem + warp * 4096;            // per-warp 4 KB landing zone, rewritten every step
    const int *rw = rows + ((size_t)blockIdx.x * nw + warp) * steps * 32;
    uint32_t acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int st = 0; st < steps; ++st) {
        const int myrow = __ldg(rw + st * 32 + lane);
        if (mode == 0) {
            const uint4 *src = tab + (size_t)myrow * 8;
            uint4 v[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) v[c] = __ldg(src + c);
#pragma unroll
            for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4 *>(buf + c * 512 + lane * 16) = v[c];
        } else if (mode == 1 || mode == 4 || mode == 3) {
            uint4 v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {        // instruction i: rows 8 (i / 2) .. +7, chunks 4 (i & 1) .. +3
                const int rr = 8 * (i >> 1) + (lane & 7), c = 4 * (i & 1) + (lane >> 3);
                const int row = __shfl_sync(0xffffffffu, myrow, rr);
                const uint4 *src = tab + (size_t)row * 8 + c;
                const uint32_t dst = ptx::smem_u32(buf + c * 512 + rr * 16);
                if (mode == 1) ptx::cp_async16_cg(dst, src, 16);
                else if (mode == 4) ptx::cp_async16_ca(dst, src, 16);
                else v[i] = __ldg(src);
            }
            if (mode == 3) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int rr = 8 * (i >> 1) + (lane & 7), c = 4 * (i & 1) + (lane >> 3);
                    *reinterpret_cast<uint4 *>(buf + c * 512 + rr * 16) = v[i];
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {        // instruction i: rows 4 i .. +3, chunks 0..7
                const int rr = 4 * i + (lane >> 3), c = lane & 7;
                const int row = __shfl_sync(0xffffffffu, myrow, rr);
                ptx::cp_async16_cg(ptx::smem_u32(buf + c * 512 + rr * 16), tab + (size_t)row * 8 + c, 16);
            }
        }
        if (mode != 0 && mode != 3) {
            asm volatile("cp.async.commit_group;" ::: "memory");
            if (depth <= 1) asm volatile("cp.async.wait_group 0;" ::: "memory");
            else if (depth == 2) asm volatile("cp.async.wait_group 1;" ::: "memory");
            else if (depth <= 4) asm volatile("cp.async.wait_group 3;" ::: "memory");
            else asm volatile("cp.async.wait_group 7;" ::: "memory");
        }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();
    const long long t1 = clock64();
    acc += ((uint32_t *)buf)[lane];
    if (tid == 0) cyc[blockIdx.x] = t1 - t0;
    if (acc == 0x12345678u) sink[0] = acc;
}

int main(int argc, char **argv) {
    const int mode = argc > 1 ? atoi(argv[1]) : 1, warps = argc > 2 ? atoi(argv[2]) : 16, depth = argc > 3 ? atoi(argv[3]) : 2;
    const int steps = argc > 4 ? atoi(argv[4]) : 64;
    const int grid_arg = argc > 5 ? atoi(argv[5]) : 0, rows_arg = argc > 6 ? atoi(argv[6]) : 131072;
    int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    if (grid_arg > 0) sms = grid_arg;
    const int R = rows_arg;
    uint4 *tab; CK(cudaMalloc(&tab, (size_t)R * 128)); CK(cudaMemset(tab, 1, (size_t)R * 128));
    std::vector<int> hr((size_t)sms * warps * steps * 32);
    srand(3); for (auto &v : hr) v = rand() % R;
    if (getenv("SEQ")) for (size_t i = 0; i < hr.size(); ++i) hr[i] = (int)(i % R);    // streaming order: consecutive rows
    int *rows; CK(cudaMalloc(&rows, hr.size() * 4)); CK(cudaMemcpy(rows, hr.data(), hr.size() * 4, cudaMemcpyHostToDevice));
    long long *cyc; CK(cudaMalloc(&cyc, sms * 8)); uint32_t *sink; CK(cudaMalloc(&sink, 4));
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, warps * 4096));
    for (int rep = 0; rep < 2; ++rep) k<<<sms, warps * 32, warps * 4096>>>(tab, rows, steps, mode, depth, cyc, sink);
    CK(cudaDeviceSynchronize());
    std::vector<long long> hc(sms); CK(cudaMemcpy(hc.data(), cyc, sms * 8, cudaMemcpyDeviceToHost));
    double mean = 0; for (auto c : hc) mean += c; mean /= sms;
    printf("grid %d table %d KB mode %d warps %2d depth %d: %.0f cycles for %d B per SM = %.1f B/cycle/SM (%.0f cycles per 32 KB tile)\n", sms, R / 8, mode, warps, depth, mean,
           warps * steps * 4096, (double)warps * steps * 4096 / mean, mean / (warps * steps * 4096 / 32768.0));
    return 0;
}
