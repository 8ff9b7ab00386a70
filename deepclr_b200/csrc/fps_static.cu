// dclr_fps_static: farthest point sampling for callers that may not allocate (CUDA graph capture, serving loops, harnesses
// that forbid cudaMalloc).  Same kernels and results as dclr_fps; the one piece of scratch that entry takes from the
// stream-ordered pool -- the Z-ordered records of the bucketed kernels (fps_sort_kernel -> fps_tmem_kernel) -- lives
// here in a file-scope device array of fixed size instead.  No counterpart in the
// reference.
//
// The array is part of the library image: every GPU that loads libdeepclr_b200.so holds its 2 MiB until the process
// exits.  All calls share it, whatever their stream, so two calls that overlap on the GPU race on it; calls issued to
// one stream are ordered and safe.  Nothing in it has to start at zero: fps_sort_kernel rewrites records [0, B*N) of
// every call before the sampling kernel reads exactly those, so the previous call's contents are never seen.
#include "common.cuh"
#include "kernels.cuh"

#include <atomic>

namespace dclr {

namespace {

constexpr int FPS_STATIC_MAX_N = 16384;                 // per cloud: the shared-memory-resident bucketed kernels' limit
constexpr int FPS_STATIC_MAX_RECORDS = 8 * 16384;       // B * N over one call: 131072 records x 16 B = 2 MiB

static __device__ __align__(256) float4 g_fps_static_records[FPS_STATIC_MAX_RECORDS];

// In host code `g_fps_static_records` is only a handle for the runtime, not a device address: the address the kernels
// get is asked for once per device (with that device current) and kept.
void *g_records_ptr[64];
std::atomic<int> g_records_state[64];                   // 0 unknown, 1 being resolved, 2 known

int static_records(void **out) {
    int dev = 0;
    DCLR_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(DCLR_E_UNSUPPORTED, "device ordinal %d not supported", dev);
    if (g_records_state[dev].load(std::memory_order_acquire) != 2) {
        int expect = 0;
        if (g_records_state[dev].compare_exchange_strong(expect, 1, std::memory_order_acquire)) {
            void *p = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&p, g_fps_static_records);
            if (e != cudaSuccess || !p) {
                g_records_state[dev].store(0, std::memory_order_release);
                return fail(DCLR_E_CUDA, "cudaGetSymbolAddress(g_fps_static_records) on device %d failed: %s", dev,
                            cudaGetErrorString(e));
            }
            g_records_ptr[dev] = p;
            g_records_state[dev].store(2, std::memory_order_release);
        } else {
            while (g_records_state[dev].load(std::memory_order_acquire) == 1) {}
            if (g_records_state[dev].load(std::memory_order_acquire) != 2)
                return fail(DCLR_E_CUDA, "dclr_fps_static: the scratch address of device %d could not be resolved", dev);
        }
    }
    *out = g_records_ptr[dev];
    return DCLR_OK;
}

}  // namespace

}  // namespace dclr

extern "C" int dclr_fps_static_limits(int *max_points_per_cloud, int *max_points_per_call) {
    if (max_points_per_cloud) *max_points_per_cloud = dclr::FPS_STATIC_MAX_N;
    if (max_points_per_call) *max_points_per_call = dclr::FPS_STATIC_MAX_RECORDS;
    return DCLR_OK;
}

extern "C" int dclr_fps_static(const float *pts, int pt_stride, int B, int N, int npoint, int tie_rule, int32_t *idx_out,
                               float *new_xyz_out, void *stream) {
    using namespace dclr;
    DCLR_CHECK_ARG(pts && idx_out, "dclr_fps_static: null pointer");
    DCLR_CHECK_ARG(tie_rule == DCLR_TIE_LOWEST_INDEX || tie_rule == DCLR_TIE_REFERENCE,
                   "dclr_fps_static: tie_rule must be 0 (lowest index) or 1 (reference kernel order), got %d", tie_rule);
    DCLR_CHECK_ARG(pt_stride >= 3 && B >= 0 && N >= 1 && npoint >= 0,
                   "dclr_fps_static: bad shape B=%d N=%d npoint=%d stride=%d", B, N, npoint, pt_stride);
    DCLR_CHECK_ARG(N <= FPS_STATIC_MAX_N, "dclr_fps_static: N=%d points per cloud is beyond the static scratch limit of %d "
                   "(use dclr_fps)", N, FPS_STATIC_MAX_N);
    DCLR_CHECK_ARG((long long)B * N <= FPS_STATIC_MAX_RECORDS, "dclr_fps_static: B*N=%lld points per call is beyond the "
                   "static scratch limit of %d (split the batch or use dclr_fps)", (long long)B * N, FPS_STATIC_MAX_RECORDS);
    if (B == 0 || npoint == 0) return DCLR_OK;
    cudaStream_t st = (cudaStream_t)stream;
    void *records = nullptr;
    { int rc_ = static_records(&records); if (rc_ != DCLR_OK) return rc_; }
    int used = 0;
    int rc = fps_bucket_launch(pts, pt_stride, B, N, npoint, tie_rule, idx_out, new_xyz_out, records, st, &used);
    if (rc != DCLR_OK || used) return rc;
    // shapes the bucketed kernels leave alone (N <= 8192, npoint < 2): register/shared-memory resident kernel, no scratch
    return fps_launch(pts, pt_stride, B, N, npoint, tie_rule, idx_out, new_xyz_out, nullptr, st);
}
