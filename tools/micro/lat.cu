// Dependent-chain latencies on sm_100a of the primitives FPS's per-round critical path is made of.
#include <cstdio>
#include <cuda_runtime.h>
#define REP 256
__global__ void k(long long *out, int *sink, int seed) {
    __shared__ int s[1024];
    int lane = threadIdx.x & 31;
    s[threadIdx.x] = threadIdx.x ^ seed;
    __syncthreads();
    int v = seed + lane;
    long long t0, t1;
    // 1. redux max chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; ++i) v = __reduce_max_sync(0xffffffffu, v + lane) ^ i;
    t1 = clock64();
    if (threadIdx.x == 0) out[0] = (t1 - t0);
    // 2. shfl chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; ++i) v = __shfl_xor_sync(0xffffffffu, v, 1) + i;
    t1 = clock64();
    if (threadIdx.x == 0) out[1] = (t1 - t0);
    // 3. ballot chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; ++i) v = __ballot_sync(0xffffffffu, (v + lane) & 1) + i;
    t1 = clock64();
    if (threadIdx.x == 0) out[2] = (t1 - t0);
    // 4. LDS chain
    int a = (v & 1023);
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; ++i) a = s[a & 1023];
    t1 = clock64();
    if (threadIdx.x == 0) out[3] = (t1 - t0);
    v += a;
    // 5. barrier chain (all warps)
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; ++i) __syncthreads();
    t1 = clock64();
    if (threadIdx.x == 0) out[4] = (t1 - t0);
    // 6. STS -> barrier -> LDS -> redux round trip
    t0 = clock64();
    for (int i = 0; i < REP; ++i) {
        if (lane == 0) s[(i & 1) * 32 + (threadIdx.x >> 5)] = v;
        __syncthreads();
        int e = s[(i & 1) * 32 + (lane & 7)];
        v = __reduce_max_sync(0xffffffffu, e) + i;
    }
    t1 = clock64();
    if (threadIdx.x == 0) out[5] = (t1 - t0);
    // 7. packed fma chain
    float f = (float)v;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; ++i) f = fmaf(f, 1.0001f, 0.5f);
    t1 = clock64();
    if (threadIdx.x == 0) out[6] = (t1 - t0);
    // 8. redux min after redux max (two dependent)
    t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < REP; ++i) {
        int m = __reduce_max_sync(0xffffffffu, v + lane);
        unsigned key = (v + lane == m) ? (unsigned)lane + i : 0xffffffffu;
        v = (int)__reduce_min_sync(0xffffffffu, key);
    }
    t1 = clock64();
    if (threadIdx.x == 0) out[7] = (t1 - t0);
    // 9. shared atomicMax chain (one address per warp)
    t0 = clock64();
    for (int i = 0; i < REP; ++i) { if (lane == 0) v = atomicMax(&s[threadIdx.x >> 5], v + i); v = __shfl_sync(0xffffffffu, v, 0); }
    t1 = clock64();
    if (threadIdx.x == 0) out[8] = (t1 - t0);
    // 10. match_any chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; ++i) v = __match_any_sync(0xffffffffu, v & 3) + i;
    t1 = clock64();
    if (threadIdx.x == 0) out[9] = (t1 - t0);
    sink[threadIdx.x] = v + (int)f;
}
int main() {
    long long *o; int *sink;
    cudaMallocManaged(&o, 16 * sizeof(long long)); cudaMalloc(&sink, 4096);
    const char *names[] = {"redux.max", "shfl", "ballot", "LDS", "bar.sync", "STS+bar+LDS+redux", "ffma", "redux.max+sel+redux.min", "atomicMax.smem+shfl", "match_any"};
    for (int threads : {32, 256, 1024}) {
        k<<<1, threads>>>(o, sink, 3); cudaDeviceSynchronize();
        k<<<1, threads>>>(o, sink, 5); cudaDeviceSynchronize();
        printf("threads=%d:", threads);
        for (int i = 0; i < 10; ++i) printf("  %s=%.1f", names[i], (double)o[i] / REP);
        printf("\n");
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
