// Dependent-chain latencies (cycles) of what the TMEM FPS kernel's round is made of, one warp and 8 warps.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define REP 128
__device__ __forceinline__ void ld2(uint32_t a, uint32_t &x, uint32_t &y) { asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(x), "=r"(y) : "r"(a) : "memory"); }
__device__ __forceinline__ void st2(uint32_t a, uint32_t x, uint32_t y) { asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(a), "r"(x), "r"(y) : "memory"); }
__device__ __forceinline__ void wld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void wst() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float fmax3(float a, float b, float c) { float r; asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) { uint64_t r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__global__ void k(long long *out, int *sink, int seed) {
    __shared__ uint32_t s_t;
    __shared__ int s[1024];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_t)), "r"(128) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    s[threadIdx.x] = threadIdx.x ^ seed;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t t = s_t + ((uint32_t)(32 * (warp & 3)) << 16) + (warp >> 2) * 64;
    for (int i = 0; i < 32; ++i) st2(t + 2 * i, i + lane, i);
    wst();
    long long t0, t1;
    uint32_t a = seed & 1, b = 0;
    int v = seed + lane;
    // 0: ld2 + wait chain (address depends on the loaded value)
    t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < REP; ++i) { ld2(t + 2 * (a & 31), a, b); wld(); a += b & 1; }
    t1 = clock64();
    if (threadIdx.x == 0) out[0] = t1 - t0;
    // 1: st2 + wait::st + ld2 + wait::ld round trip
    t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < REP; ++i) { st2(t + 2 * (a & 31), a + 1, b); wst(); ld2(t + 2 * (a & 31), a, b); wld(); }
    t1 = clock64();
    if (threadIdx.x == 0) out[1] = t1 - t0;
    // 2: ffs chain
    v = (int)a | 0x100;
    t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < REP; ++i) v = (v >> (__ffs(v) & 1)) | 0x10000 | i;
    t1 = clock64();
    if (threadIdx.x == 0) out[2] = t1 - t0;
    // 3: fmax3 chain
    float f = (float)v;
    t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < REP; ++i) f = fmax3(f - 1.0f, 1.0f - f, 0.f);
    t1 = clock64();
    if (threadIdx.x == 0) out[3] = t1 - t0;
    // 4: packed sub + fma chain (2 dependent packed ops)
    uint64_t p = ((uint64_t)__float_as_uint(f) << 32) | (uint32_t)v;
    t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < REP; ++i) { p = sub2(p, 0x3f8000003f800000ull); p = fma2(p, p, p); }
    t1 = clock64();
    if (threadIdx.x == 0) out[4] = t1 - t0;
    // 5: redux -> compare -> select -> redux (the key chain) + ballot in parallel
    v += (int)(p >> 40);
    t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < REP; ++i) {
        const int m = __reduce_max_sync(0xffffffffu, v + lane);
        const unsigned key = (v + lane == m) ? (unsigned)lane + i : 0xffffffffu;
        const unsigned hb = __ballot_sync(0xffffffffu, v + lane == m);
        const int m2 = __reduce_max_sync(0xffffffffu, v + lane == m ? 0 : v + lane);
        v = (int)__reduce_min_sync(0xffffffffu, key) + __popc(hb) + (m2 & 1);
    }
    t1 = clock64();
    if (threadIdx.x == 0) out[5] = t1 - t0;
    // 6: LDS -> packed math -> fmin chain as in the update (address from previous value)
    int ad = v & 1023;
    t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < REP; ++i) { ad = (s[ad & 1023] + i) & 1023; }
    t1 = clock64();
    if (threadIdx.x == 0) out[6] = t1 - t0;
    // 7: taken-branch chain (uniform, data dependent)
    v += ad;
    t0 = clock64();
    for (int i = 0; i < REP; ++i) {
        if (v & 1) v = v * 3 + 1; else v = v >> 1;
        if (v & 2) v += 5; else v ^= 9;
    }
    t1 = clock64();
    if (threadIdx.x == 0) out[7] = t1 - t0;
    // 8: clock64 back to back
    t0 = clock64();
    long long acc = 0;
#pragma unroll 8
    for (int i = 0; i < REP; ++i) acc += clock64();
    t1 = clock64();
    if (threadIdx.x == 0) out[8] = t1 - t0;
    sink[threadIdx.x] = v + (int)f + (int)acc + (int)p + a;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(s_t), "r"(128) : "memory");
}
int main() {
    long long *o; int *sink;
    cudaMallocManaged(&o, 16 * sizeof(long long)); cudaMalloc(&sink, 4096);
    const char *names[] = {"ldtm.x2+wait", "sttm+wait+ldtm+wait", "ffs", "fmax3(sub,sub)", "sub2+fma2", "redux-key-chain", "LDS+add", "2 branches", "clock64"};
    for (int threads : {32, 256}) {
        k<<<1, threads>>>(o, sink, 3); cudaDeviceSynchronize();
        k<<<1, threads>>>(o, sink, 5); cudaDeviceSynchronize();
        printf("threads=%d:", threads);
        for (int i = 0; i < 9; ++i) printf("  %s=%.1f", names[i], (double)o[i] / REP);
        printf("\n");
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
