// Error plumbing and identification for libdeepclr_b200.
#include "common.cuh"
#include <string.h>
#include <atomic>

namespace dclr {

char *err_buf() {
    static thread_local char buf[512] = {0};
    return buf;
}

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err_buf(), 512, fmt, ap);
    va_end(ap);
    return code;
}

// Scratch memory of the op-level entry points (dclr_fps, dclr_ball_query ...): a private stream-ordered pool per
// device that keeps what it has been given.  The device's DEFAULT pool has a release threshold of 0, i.e. it hands
// freed memory back to the driver at every synchronisation -- a caller that synchronises per op (the reference's
// training loop does) then pays a real cudaMalloc (~0.25 ms) inside every call.
static cudaMemPool_t g_pool[64] = {nullptr};
static std::atomic<int> g_pool_lock{0};
int scratch_alloc(void **p, size_t bytes, cudaStream_t st) {
    int dev = 0;
    DCLR_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(DCLR_E_UNSUPPORTED, "device ordinal %d not supported", dev);
    if (!g_pool[dev]) {
        while (g_pool_lock.exchange(1, std::memory_order_acquire)) {}
        int rc = DCLR_OK;
        if (!g_pool[dev]) {
            cudaMemPoolProps props = {};
            props.allocType = cudaMemAllocationTypePinned;
            props.location.type = cudaMemLocationTypeDevice;
            props.location.id = dev;
            cudaMemPool_t pool = nullptr;
            cudaError_t e = cudaMemPoolCreate(&pool, &props);
            uint64_t keep = UINT64_MAX;
            if (e == cudaSuccess) e = cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
            if (e != cudaSuccess) rc = fail(DCLR_E_CUDA, "scratch pool: %s", cudaGetErrorString(e));
            else g_pool[dev] = pool;
        }
        g_pool_lock.store(0, std::memory_order_release);
        if (rc != DCLR_OK) return rc;
    }
    DCLR_CUDA(cudaMallocFromPoolAsync(p, bytes, g_pool[dev], st));
    return DCLR_OK;
}
void scratch_free(void *p, cudaStream_t st) {
    if (p) cudaFreeAsync(p, st);
}

}  // namespace dclr

extern "C" const char *dclr_version(void) { return "deepclr_b200 0.2 (sm_100a)"; }
extern "C" const char *dclr_last_error(void) { return dclr::err_buf(); }
