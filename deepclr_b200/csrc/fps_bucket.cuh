// Helpers shared by the bucketed FPS kernels (fps_bucket.cu: registers, one sample per round; fps_tmem.cu: running minima
// in tensor memory, several samples per round).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dclr {

// ---- packed fp32x2 helpers (FADD2 / FMUL2 / FFMA2: each half rounds exactly like the scalar op)
__device__ __forceinline__ uint64_t fb_pack2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void fb_unpack2(uint64_t v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fb_sub2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t fb_mul2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t fb_fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ uint64_t fb_lds64(uint32_t addr) {
    uint64_t r;
    asm volatile("ld.shared.b64 %0, [%1];" : "=l"(r) : "r"(addr));
    return r;
}
__device__ __forceinline__ uint32_t fb_lds32(uint32_t addr) {
    uint32_t r;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(r) : "r"(addr));
    return r;
}

// Tie order of the reference kernel for N > 1024 (1024-thread block, sampling_gpu.cu:113-203): thread tid = k & 1023
// keeps its first strict maximum, the shared-memory tree keeps the LOWER slot on equal values at every level, so among
// equal maxima the slot with the smallest bit-reversed tid wins, then the smallest k.  14 bits for N <= 16384.
__device__ __forceinline__ uint32_t fb_rank16(int k) { return ((__brev((uint32_t)k & 1023u) >> 22) << 4) | ((uint32_t)k >> 10); }
__device__ __forceinline__ int fb_unrank16(uint32_t r) { return (int)((__brev(r >> 4) >> 22) | ((r & 15u) << 10)); }




}  // namespace dclr
