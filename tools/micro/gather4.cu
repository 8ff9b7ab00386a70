// TMA tile::gather4 probe: semantics (which tensor-map box shape the instruction wants, where the 4 rows land, the
// 128-byte swizzle pattern) and throughput of row gathers from an L2-resident table, all SMs at once.
//   ./gather4 <box_rows: 1|4> <swizzle: 0|1> <ops per tile> <iters> [grid]
// Table: R rows x 64 u16 (128 B per row), value(row, col) = (row * 7 + col) & 0xffff.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../../deepclr_b200/csrc/ptx.cuh"
using namespace dclr;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); fflush(stdout); exit(1); } } while (0)

__device__ __forceinline__ void gather4(uint32_t dst, const CUtensorMap *map, uint64_t *bar, int col, int r0, int r1, int r2, int r3) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(dst), "l"(map), "r"(ptx::smem_u32(bar)),
        "r"(col), "r"(r0), "r"(r1), "r"(r2), "r"(r3) : "memory");
}

__global__ void __launch_bounds__(128, 1)
k(const __grid_constant__ CUtensorMap map, const int *rows, int R, int ops, int iters, uint16_t *dump, long long *cyc) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    const int tid = threadIdx.x;
    if (tid == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
    for (int i = tid; i < ops * 512 / 4; i += blockDim.x) ((uint32_t *)smem)[i] = 0xdeadbeefu;
    ptx::fence_proxy_async_smem();
    __syncthreads();
    long long t0 = 0, t1 = 0;
    if (tid < 32) {
        uint32_t ph = 0;
        t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            const int *rw = rows + ((size_t)(blockIdx.x * iters + it) * ops) * 4;
            if (tid == 0) ptx::mbar_expect_tx(&bar, ops * 512);
            __syncwarp();
            for (int o = tid; o < ops; o += 32) {
                const int4 r = *reinterpret_cast<const int4 *>(rw + o * 4);
                gather4(ptx::smem_u32(smem + o * 512), &map, &bar, 0, r.x, r.y, r.z, r.w);
            }
            ptx::mbar_wait(&bar, ph); ph ^= 1;
        }
        t1 = clock64();
    }
    __syncthreads();
    if (blockIdx.x == 0) {
        for (int i = tid; i < ops * 256; i += blockDim.x) dump[i] = ((uint16_t *)smem)[i];
        if (tid == 0) cyc[0] = t1 - t0;
    }
}

int main(int argc, char **argv) {
    const int box_rows = argc > 1 ? atoi(argv[1]) : 4, swz = argc > 2 ? atoi(argv[2]) : 1;
    const int ops = argc > 3 ? atoi(argv[3]) : 32, iters = argc > 4 ? atoi(argv[4]) : 64;
    int dev_sms = 0;
    CK(cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, 0));
    const int grid = argc > 5 ? atoi(argv[5]) : dev_sms;
    const int R = 131072;
    std::vector<uint16_t> h((size_t)R * 64);
    for (int r = 0; r < R; ++r) for (int c = 0; c < 64; ++c) h[(size_t)r * 64 + c] = (uint16_t)((r * 7 + c) & 0xffff);
    uint16_t *d_tab; CK(cudaMalloc(&d_tab, h.size() * 2)); CK(cudaMemcpy(d_tab, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
    std::vector<int> hr((size_t)grid * iters * ops * 4);
    srand(1);
    for (auto &v : hr) v = rand() % R;
    int *d_rows; CK(cudaMalloc(&d_rows, hr.size() * 4)); CK(cudaMemcpy(d_rows, hr.data(), hr.size() * 4, cudaMemcpyHostToDevice));
    uint16_t *d_dump; CK(cudaMalloc(&d_dump, ops * 512)); long long *d_cyc; CK(cudaMalloc(&d_cyc, 8));

    typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void *fn = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    if (!fn) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
    CUtensorMap map;
    cuuint64_t gdim[2] = {64, (cuuint64_t)R}, gstr[1] = {128};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows}, es[2] = {1, 1};
    CUresult cr = ((EncodeFn)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, d_tab, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                 swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("box_rows %d swizzle %d ops %d iters %d grid %d: encode rc %d\n", box_rows, swz, ops, iters, grid, (int)cr); fflush(stdout);
    if (cr != CUDA_SUCCESS) return 1;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, ops * 512 + 1024));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k<<<grid, 128, ops * 512 + 1024>>>(map, d_rows, R, ops, 2, d_dump, d_cyc);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    k<<<grid, 128, ops * 512 + 1024>>>(map, d_rows, R, ops, iters, d_dump, d_cyc);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    long long cyc; CK(cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost));
    std::vector<uint16_t> dump(ops * 256); CK(cudaMemcpy(dump.data(), d_dump, ops * 512, cudaMemcpyDeviceToHost));
    // check the last iteration of CTA 0: op o, row j of the op at byte offset o*512 + j*128, 16-byte chunk c at c ^ (row_in_tile & 7)
    const int *rw = hr.data() + ((size_t)(0 * iters + iters - 1) * ops) * 4;
    int bad_plain = 0, bad_swz = 0;
    for (int o = 0; o < ops; ++o) for (int j = 0; j < 4; ++j) for (int c = 0; c < 64; ++c) {
        const uint16_t want = (uint16_t)((rw[o * 4 + j] * 7 + c) & 0xffff);
        const int trow = o * 4 + j;
        bad_plain += dump[trow * 64 + c] != want;
        const int chunk = (c >> 3) ^ (trow & 7);
        bad_swz += dump[trow * 64 + chunk * 8 + (c & 7)] != want;
    }
    printf("  mismatches: plain layout %d, 128B-swizzled layout %d (of %d)\n", bad_plain, bad_swz, ops * 256);
    printf("  first 8 u16 of smem: "); for (int i = 0; i < 8; ++i) printf("%04x ", dump[i]); printf("| want row %d: %04x..\n", rw[0], (rw[0] * 7) & 0xffff);
    printf("  %.1f cycles per tile of %d B (CTA 0) = %.1f B/cycle/SM; whole grid %.3f ms = %.1f GB/s\n", (double)cyc / iters, ops * 512,
           (double)ops * 512 * iters / cyc, ms, (double)grid * iters * ops * 512 / ms / 1e6);
    return 0;
}
