// Dependent-chain latency of packed fp32x2 ops, LDS.64, and a data-dependent switch (BRX) on sm_100a.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define REP 256
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) {
    uint64_t r; asm volatile("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__global__ void k(long long *out, uint64_t *sink, int seed) {
    __shared__ uint64_t s[512];
    s[threadIdx.x & 511] = (threadIdx.x * 8) & 4095 ;
    __syncthreads();
    long long t0, t1;
    uint64_t v = seed + threadIdx.x, c = 0x3f8000003f800000ull;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; ++i) v = fma2(v, c, c);
    t1 = clock64();
    if (threadIdx.x == 0) out[0] = t1 - t0;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; ++i) v = sub2(v, c);
    t1 = clock64();
    if (threadIdx.x == 0) out[1] = t1 - t0;
    // LDS.64 pointer chase
    uint64_t a = v & 511;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; ++i) a = s[(a >> 3) & 511];
    t1 = clock64();
    if (threadIdx.x == 0) out[2] = t1 - t0;
    // data dependent switch, warp-uniform
    int x = (int)(a & 31) ^ seed;
    float f0 = 1, f1 = 2, f2 = 3, f3 = 4, f4 = 5, f5 = 6, f6 = 7, f7 = 8, d = (float)seed;
    t0 = clock64();
    for (int i = 0; i < REP; ++i) {
        switch (x & 7) {
            case 0: f0 = d = fminf(d, f0) + 1; break; case 1: f1 = d = fminf(d, f1) + 2; break;
            case 2: f2 = d = fminf(d, f2) + 3; break; case 3: f3 = d = fminf(d, f3) + 4; break;
            case 4: f4 = d = fminf(d, f4) + 5; break; case 5: f5 = d = fminf(d, f5) + 6; break;
            case 6: f6 = d = fminf(d, f6) + 7; break; default: f7 = d = fminf(d, f7) + 8; break;
        }
        x = x * 5 + 1 + (int)d % 3;
    }
    t1 = clock64();
    if (threadIdx.x == 0) out[3] = t1 - t0;
    // same without the switch (arith only)
    t0 = clock64();
    for (int i = 0; i < REP; ++i) { d = fminf(d, f0) + 1; x = x * 5 + 1 + (int)d % 3; }
    t1 = clock64();
    if (threadIdx.x == 0) out[4] = t1 - t0;
    sink[threadIdx.x] = v + a + x + (uint64_t)(f0 + f1 + f2 + f3 + f4 + f5 + f6 + f7 + d);
}
int main() {
    long long *o; uint64_t *sink;
    cudaMallocManaged(&o, 64); cudaMalloc(&sink, 8192);
    for (int th : {32, 256}) {
        k<<<1, th>>>(o, sink, 3); cudaDeviceSynchronize();
        k<<<1, th>>>(o, sink, 4); cudaDeviceSynchronize();
        printf("threads=%d: fma.f32x2=%.1f sub.f32x2=%.1f LDS64chase=%.1f switch8+arith=%.1f arith-only=%.1f  [%s]\n", th,
               (double)o[0] / REP, (double)o[1] / REP, (double)o[2] / REP, (double)o[3] / REP, (double)o[4] / REP,
               cudaGetErrorString(cudaGetLastError()));
    }
}
