// Shared helpers for libdeepclr_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/deepclr_b200.h"

namespace dclr {

// thread-local error text behind dclr_last_error()
char *err_buf();
int fail(int code, const char *fmt, ...);

#define DCLR_CHECK_ARG(cond, ...)                                   \
    do {                                                            \
        if (!(cond)) return ::dclr::fail(DCLR_E_INVALID, __VA_ARGS__); \
    } while (0)

#define DCLR_CUDA(expr)                                                                                  \
    do {                                                                                                 \
        cudaError_t e__ = (expr);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return ::dclr::fail(DCLR_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, \
                                __LINE__);                                                               \
    } while (0)

#define DCLR_LAUNCH_CHECK(name)                                                                       \
    do {                                                                                              \
        cudaError_t e__ = cudaGetLastError();                                                         \
        if (e__ != cudaSuccess)                                                                       \
            return ::dclr::fail(DCLR_E_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(e__)); \
    } while (0)

// stream-ordered scratch memory for the op-level entry points (core.cu)
int scratch_alloc(void **p, size_t bytes, cudaStream_t st);
void scratch_free(void *p, cudaStream_t st);

static inline int div_up(int a, int b) { return (a + b - 1) / b; }

// The reference kernels' squared distance as nvcc contracts it (FMUL dy*dy, FFMA dx, FFMA dz): SURVEY.md App. A.
__device__ __forceinline__ float sqdist_pn2(float dx, float dy, float dz) {
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}
// torch-cluster's accumulation order: 0 + dx*dx, then dy, then dz.
__device__ __forceinline__ float sqdist_tc(float dx, float dy, float dz) {
    return __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
}

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }

}  // namespace dclr
