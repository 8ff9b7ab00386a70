// Farthest point sampling for sm_100a: one CTA per cloud, the cloud held in shared memory (SoA) and, per thread,
// in registers (as packed fp32x2 pairs); min-distance state never leaves registers; argmax = redux.sync within
// the warp, one shared-memory slot per warp, ONE __syncthreads per round.
//
// Replaces furthest_point_sampling_kernel + gather_points_kernel_fast of the reference
// (extern/Pointnet2.PyTorch/pointnet2/src/sampling_gpu.cu:94-209, :8-24).  Arithmetic per round and point k is the
// reference's (sampling_gpu.cu:133-137 as contracted by nvcc):
//     d  = fma(dz,dz, fma(dx,dx, dy*dy)),  d* = p_k - p_old
//     t_k = min(d, t_k)           (t_k starts at 1e10: pointnet2_utils.py:26)
//     next = argmax_k t_k         ties -> LOWEST point index (BASELINE.json's stated rule)
#include "common.cuh"
#include "kernels.cuh"
#include "ptx.cuh"
#include <limits.h>

namespace dclr {

constexpr int FPS_MAX_RESIDENT_N = 16384;

// Packed fp32x2 arithmetic (sm_100 FADD2 / FMUL2 / FFMA2): two points per instruction, each half rounded exactly like
// the scalar IEEE op, so the result is bit-identical to the scalar expression above.
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ float max3(float a, float b, float c) {
    float r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}

__device__ __forceinline__ uint64_t lds64(uint32_t addr) {
    uint64_t r;
    asm volatile("ld.shared.b64 %0, [%1];" : "=l"(r) : "r"(addr));
    return r;
}

// T threads, PP point PAIRS per thread: thread tid owns points 2*(tid + i*T) and 2*(tid + i*T) + 1, i < PP.  The
// shared arrays are padded to the full capacity CAP = 2*T*PP so that no load needs a bounds guard (the padding never
// wins: its running minimum starts at -1).  ZS: read z from shared memory (one 8-byte load per pair, immediate
// offsets) instead of registers -- needed at T=1024, PP=8 where only 64 registers per thread exist.
// Per round: packed distance update -> warp redux max -> slot per warp -> barrier -> block max; only the thread(s)
// holding the block max look up their lowest index and atomicMin it -> barrier.
// Tie order.  tieL < 0: lowest point index (the stated default).  tieL >= 0: the reference kernel's own order
// (sampling_gpu.cu:113-203 with a block of 2^tieL threads = opt_n_threads(N), cuda_utils.h:10-14): thread k mod 2^tieL
// keeps its FIRST strict maximum, and every level of the shared-memory tree keeps the LOWER slot on equal values, so
// among equal maxima the smallest bit-reversed slot wins, then the smallest k.  rank() is that order as an integer.
__device__ __forceinline__ int tie_rank(int k, int L) {
    if (L <= 0) return k;
    return (int)(((__brev((uint32_t)k & ((1u << L) - 1u)) >> (32 - L)) << (31 - L)) | ((uint32_t)k >> L));
}
__device__ __forceinline__ int tie_unrank(int r, int L) {
    if (L <= 0) return r;
    const uint32_t slot = __brev((uint32_t)r >> (31 - L)) >> (32 - L), q = (uint32_t)r & ((1u << (31 - L)) - 1u);
    return (int)((q << L) | slot);
}

template <int T, int PP, bool ZS>
__global__ void __launch_bounds__(T, 1)
fps_resident_kernel(const float *__restrict__ pts, int stride, int N, int npoint, int tieL,
                    int32_t *__restrict__ idx_out, float *__restrict__ new_xyz_out) {
    extern __shared__ __align__(16) float smem[];
    constexpr int CAP = 2 * T * PP;
    constexpr int NW = T / 32;
    float *s_x = smem, *s_y = smem + CAP, *s_z = smem + 2 * CAP;
    int *s_wmax = reinterpret_cast<int *>(smem + 3 * CAP);          // [2][32]
    int *s_win = s_wmax + 64;                                       // [2]

    const int b = blockIdx.x;
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    pts += (size_t)b * N * stride;
    idx_out += (size_t)b * npoint;
    if (new_xyz_out) new_xyz_out += (size_t)b * npoint * 3;

    // stage the cloud: coalesced (float4 when the point stride is 4 floats); zero the padding
    if (stride == 4) {
        const float4 *p4 = reinterpret_cast<const float4 *>(pts);
        for (int k = tid; k < N; k += T) {
            float4 v = __ldg(p4 + k);
            s_x[k] = v.x; s_y[k] = v.y; s_z[k] = v.z;
        }
    } else {
        for (int k = tid; k < N; k += T) {
            const float *p = pts + (size_t)k * stride;
            s_x[k] = __ldg(p); s_y[k] = __ldg(p + 1); s_z[k] = __ldg(p + 2);
        }
    }
    for (int k = N + tid; k < CAP; k += T) { s_x[k] = 0.f; s_y[k] = 0.f; s_z[k] = 0.f; }
    if (tid < 2) s_win[tid] = INT_MAX;
    __syncthreads();

    uint64_t x2[PP], y2[PP], z2[ZS ? 1 : PP];
    float tl[PP], th[PP];
    const uint32_t zaddr = ptx::smem_u32(s_z) + tid * 8;
#pragma unroll
    for (int i = 0; i < PP; ++i) {
        const int k = 2 * (tid + i * T);
        x2[i] = *reinterpret_cast<const uint64_t *>(s_x + k);
        y2[i] = *reinterpret_cast<const uint64_t *>(s_y + k);
        if (!ZS) z2[i] = *reinterpret_cast<const uint64_t *>(s_z + k);
        tl[i] = k < N ? 1e10f : -1.0f;        // -1 can never win (reference: best starts at -1, strict >)
        th[i] = k + 1 < N ? 1e10f : -1.0f;
    }

    int old = 0;
    if (tid == 0) idx_out[0] = 0;
    for (int j = 1; j < npoint; ++j) {
        const float x1 = s_x[old], y1 = s_y[old], z1 = s_z[old];
        if (tid == 0 && new_xyz_out) {
            new_xyz_out[(j - 1) * 3 + 0] = x1; new_xyz_out[(j - 1) * 3 + 1] = y1; new_xyz_out[(j - 1) * 3 + 2] = z1;
        }
        const uint64_t xo = pack2(x1, x1), yo = pack2(y1, y1), zo = pack2(z1, z1);
        float best = -1.0f;
#pragma unroll
        for (int i = 0; i < PP; ++i) {
            const uint64_t zz = ZS ? lds64(zaddr + i * (T * 8)) : z2[i];
            const uint64_t dx = sub2(x2[i], xo), dy = sub2(y2[i], yo), dz = sub2(zz, zo);
            const uint64_t d2 = fma2(dz, dz, fma2(dx, dx, mul2(dy, dy)));
            float dl, dh;
            unpack2(d2, dl, dh);
            tl[i] = fminf(dl, tl[i]);
            th[i] = fminf(dh, th[i]);
            best = max3(best, tl[i], th[i]);
        }
        // block max of the float through its integer image (values are >= 0 or exactly -1.0f, a negative int)
        const int bi = __float_as_int(best);
        const int wm = __reduce_max_sync(0xffffffffu, bi);
        const int par = j & 1;
        if (lane == 0) s_wmax[par * 32 + warp] = wm;
        __syncthreads();
        const int m = __reduce_max_sync(0xffffffffu, lane < NW ? s_wmax[par * 32 + lane] : INT_MIN);
        if (tid == 0) s_win[par ^ 1] = INT_MAX;                  // re-arm the other slot for the next round
        if (bi == m) {
            int cand = INT_MAX;
#pragma unroll
            for (int i = 0; i < PP; ++i) {                       // lowest rank (= lowest index when tieL < 0) survives
                const int k = 2 * (tid + i * T);
                if (__float_as_int(tl[i]) == m) cand = min(cand, tie_rank(k, tieL));
                if (__float_as_int(th[i]) == m) cand = min(cand, tie_rank(k + 1, tieL));
            }
            atomicMin(&s_win[par], cand);
        }
        __syncthreads();
        old = tie_unrank(s_win[par], tieL);
        if (tid == 0) idx_out[j] = old;
    }
    if (tid == 0 && new_xyz_out && npoint > 0) {
        new_xyz_out[(npoint - 1) * 3 + 0] = s_x[old];
        new_xyz_out[(npoint - 1) * 3 + 1] = s_y[old];
        new_xyz_out[(npoint - 1) * 3 + 2] = s_z[old];
    }
}

// Clouds too large for one SM's shared memory: state streamed from global/L2 each round (same arithmetic and
// tie-break).  temp: [B,N] scratch.
template <int T>
__global__ void __launch_bounds__(T, 1)
fps_stream_kernel(const float *__restrict__ pts, int stride, int N, int npoint, int tieL, float *__restrict__ temp,
                  int32_t *__restrict__ idx_out, float *__restrict__ new_xyz_out) {
    static_assert(T == 1024, "thread k mod T must own the points the reference's 1024-thread block gives it");
    __shared__ int s_wmax[32];
    __shared__ int s_win[2];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pts += (size_t)b * N * stride;
    temp += (size_t)b * N;
    idx_out += (size_t)b * npoint;
    if (new_xyz_out) new_xyz_out += (size_t)b * npoint * 3;
    for (int k = tid; k < N; k += T) temp[k] = 1e10f;
    if (tid < 2) s_win[tid] = INT_MAX;
    __syncthreads();
    int old = 0;
    if (tid == 0) idx_out[0] = 0;
    for (int j = 1; j <= npoint; ++j) {
        const float x1 = __ldg(pts + (size_t)old * stride), y1 = __ldg(pts + (size_t)old * stride + 1),
                    z1 = __ldg(pts + (size_t)old * stride + 2);
        if (tid == 0 && new_xyz_out) {
            new_xyz_out[(j - 1) * 3 + 0] = x1; new_xyz_out[(j - 1) * 3 + 1] = y1; new_xyz_out[(j - 1) * 3 + 2] = z1;
        }
        if (j == npoint) break;
        float best = -1.0f;
        int besti = INT_MAX;
        for (int k = tid; k < N; k += T) {
            const float *p = pts + (size_t)k * stride;
            float d = sqdist_pn2(__ldg(p) - x1, __ldg(p + 1) - y1, __ldg(p + 2) - z1);
            float d2 = fminf(d, temp[k]);
            temp[k] = d2;
            if (d2 > best) { best = d2; besti = k; }      // strict >: lowest k of this thread
        }
        int wb = __reduce_max_sync(0xffffffffu, __float_as_int(best));
        if (lane == 0) s_wmax[warp] = wb;
        __syncthreads();
        int v = lane < (T / 32) ? s_wmax[lane] : INT_MIN;
        const int m = __reduce_max_sync(0xffffffffu, v);
        if (tid == 0) s_win[(j + 1) & 1] = INT_MAX;
        if (__float_as_int(best) == m) atomicMin(&s_win[j & 1], tie_rank(besti, tieL));
        __syncthreads();
        old = tie_unrank(s_win[j & 1], tieL);
        if (tid == 0) idx_out[j] = old;
    }
}

template <int T, int PP, bool ZS>
static int launch_resident(const float *pts, int stride, int B, int N, int npoint, int tieL, int32_t *idx, float *nx,
                           cudaStream_t st) {
    size_t smem = (size_t)3 * (2 * T * PP) * sizeof(float) + 66 * sizeof(int);
    auto kern = fps_resident_kernel<T, PP, ZS>;
    DCLR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<B, T, smem, st>>>(pts, stride, N, npoint, tieL, idx, nx);
    DCLR_LAUNCH_CHECK("fps_resident_kernel");
    return DCLR_OK;
}

int fps_launch(const float *pts, int stride, int B, int N, int npoint, int tie_rule, int32_t *idx, float *nx,
               float *temp_big, cudaStream_t st) {
    int tieL = -1;
    if (tie_rule == DCLR_TIE_REFERENCE)
        for (tieL = 0; (2 << tieL) <= N && tieL < 10; ++tieL) {}          // log2(opt_n_threads(N)), cuda_utils.h:10-14
    // T threads x PP pairs x 2 points >= N
    if (N <= 256) return launch_resident<128, 1, false>(pts, stride, B, N, npoint, tieL, idx, nx, st);
    if (N <= 512) return launch_resident<128, 2, false>(pts, stride, B, N, npoint, tieL, idx, nx, st);
    if (N <= 1024) return launch_resident<128, 4, false>(pts, stride, B, N, npoint, tieL, idx, nx, st);
    if (N <= 2048) return launch_resident<256, 4, false>(pts, stride, B, N, npoint, tieL, idx, nx, st);
    if (N <= 4096) return launch_resident<256, 8, false>(pts, stride, B, N, npoint, tieL, idx, nx, st);
    if (N <= 8192) return launch_resident<512, 8, false>(pts, stride, B, N, npoint, tieL, idx, nx, st);
    if (N <= FPS_MAX_RESIDENT_N) return launch_resident<1024, 8, true>(pts, stride, B, N, npoint, tieL, idx, nx, st);
    if (!temp_big) return fail(DCLR_E_INVALID, "fps: N=%d needs a scratch buffer", N);
    fps_stream_kernel<1024><<<B, 1024, 0, st>>>(pts, stride, N, npoint, tieL, temp_big, idx, nx);
    DCLR_LAUNCH_CHECK("fps_stream_kernel");
    return DCLR_OK;
}

}  // namespace dclr

extern "C" int dclr_fps(const float *pts, int pt_stride, int B, int N, int npoint, int tie_rule, int32_t *idx_out,
                        float *new_xyz_out, void *stream) {
    using namespace dclr;
    DCLR_CHECK_ARG(pts && idx_out, "dclr_fps: null pointer");
    DCLR_CHECK_ARG(tie_rule == DCLR_TIE_LOWEST_INDEX || tie_rule == DCLR_TIE_REFERENCE,
                   "dclr_fps: tie_rule must be 0 (lowest index) or 1 (reference kernel order), got %d", tie_rule);
    DCLR_CHECK_ARG(pt_stride >= 3 && B >= 0 && N >= 1 && npoint >= 0, "dclr_fps: bad shape B=%d N=%d npoint=%d stride=%d",
                   B, N, npoint, pt_stride);
    if (B == 0 || npoint == 0) return DCLR_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (N > 8192 && N <= 131072) {                         // exact pruned variants (fps_bucket.cu, fps_bucket_l2.cu)
        void *ws = nullptr;
        { int rc_ = scratch_alloc(&ws, fps_bucket_workspace_bytes(B, N), st); if (rc_ != DCLR_OK) return rc_; }
        int used = 0;
        int rc = fps_bucket_launch(pts, pt_stride, B, N, npoint, tie_rule, idx_out, new_xyz_out, ws, st, &used);
        scratch_free(ws, st);
        if (rc != DCLR_OK || used) return rc;
    }
    float *temp = nullptr;
    if (N > FPS_MAX_RESIDENT_N) { int rc_ = scratch_alloc((void **)&temp, (size_t)B * N * sizeof(float), st); if (rc_ != DCLR_OK) return rc_; }
    int rc = fps_launch(pts, pt_stride, B, N, npoint, tie_rule, idx_out, new_xyz_out, temp, st);
    scratch_free(temp, st);
    return rc;
}
