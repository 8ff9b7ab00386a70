// Bucketed farthest point sampling (8192 < N <= 16384; up to 131072 with the point state in L2, fps_bucket_l2.cu): the selection of fps.cu with the N x npoint distance updates
// pruned EXACTLY.
//
// The cloud is first put into Z-order (fps_sort_kernel: 64 x 64 Morton cells over the two widest axes, counting sort
// in shared memory) so that every run of 64 consecutive records -- a BUCKET -- is spatially compact.  Per bucket the
// sampling kernel keeps the bounding box, the maximum of the running min-distances t_k and the (lowest original index,
// position) at that maximum.  A new sample lowers t_k only where d(k, new) < t_k, so a bucket is skipped when a lower
// bound of the distance from the sample to its box (shrunk by 1e-4, far more than the rounding of the per-point
// expression) is >= the bucket's max t: the update would be a no-op for each of its points.  Updated buckets use the
// reference's arithmetic (sampling_gpu.cu:133-137 as contracted: fma(dz,dz,fma(dx,dx,dy*dy))) on the same operands,
// so every t_k and every selected index is bit-identical to the unpruned kernel and to the reference.  Ties -> lowest
// ORIGINAL index, as everywhere in this library.  In steady state a round touches a handful of the 256 buckets
// instead of 16384 points; what remains per round is the block-wide argmax (one barrier).
//
// Ownership: warp w of NW owns buckets w, w+NW, ... (neighbouring buckets land on different warps); lane l owns records
// 2l, 2l+1 of each of them (t in registers, coordinates in shared memory); lane i < PP keeps the metadata of the warp's
// i-th bucket in registers, so the skip test of all of a warp's buckets is one predicated instruction sequence + ballot.
#include "common.cuh"
#include "kernels.cuh"
#include "ptx.cuh"
#include "fps_bucket.cuh"
#include <float.h>
#include <limits.h>
#include <stdlib.h>
#include <type_traits>

namespace dclr {

constexpr int FS_G = 64;                   // Morton cells per axis
constexpr int FS_CELLS = FS_G * FS_G;
constexpr int FS_T = 1024;

__device__ __forceinline__ uint32_t spread6(uint32_t v) {       // abcdef -> 0a0b0c0d0e0f
    v &= 0x3fu;
    v = (v | (v << 4)) & 0x30fu;
    v = (v | (v << 2)) & 0x333u;
    v = (v | (v << 1)) & 0x555u;
    return v;
}
__device__ __forceinline__ int fs_cell(float v0, float v1, float o0, float o1, float inv0, float inv1) {
    const int c0 = (int)fminf(fmaxf(floorf((v0 - o0) * inv0), 0.f), (float)(FS_G - 1));
    const int c1 = (int)fminf(fmaxf(floorf((v1 - o1) * inv1), 0.f), (float)(FS_G - 1));
    return (int)(spread6((uint32_t)c0) | (spread6((uint32_t)c1) << 1));
}

// one CTA per cloud: records (x, y, z, original index) in Morton-cell order
__global__ void __launch_bounds__(FS_T, 1)
fps_sort_kernel(const float *__restrict__ pts, int stride, int N, float4 *__restrict__ sorted) {
    __shared__ int s_cnt[FS_CELLS];
    __shared__ float s_red[6][32];
    __shared__ float s_par[6];
    __shared__ int s_wsum[32];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pts += (size_t)b * N * stride;
    sorted += (size_t)b * N;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int k = tid; k < N; k += FS_T) {
        const float *p = pts + (size_t)k * stride;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float v = __ldg(p + a);
            lo[a] = fminf(lo[a], v);
            hi[a] = fmaxf(hi[a], v);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], d));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], d));
        }
        if (lane == 0) { s_red[a][warp] = lo[a]; s_red[3 + a][warp] = hi[a]; }
    }
    for (int e = tid; e < FS_CELLS; e += FS_T) s_cnt[e] = 0;
    __syncthreads();
    if (tid == 0) {
        float l[3], ext[3];
        for (int a = 0; a < 3; ++a) {
            float h = -FLT_MAX;
            l[a] = FLT_MAX;
            for (int w = 0; w < FS_T / 32; ++w) { l[a] = fminf(l[a], s_red[a][w]); h = fmaxf(h, s_red[3 + a][w]); }
            ext[a] = h - l[a];
        }
        int drop = 0;
        if (ext[1] < ext[drop]) drop = 1;
        if (ext[2] < ext[drop]) drop = 2;
        const int a0 = drop == 0 ? 1 : 0, a1 = drop == 2 ? 1 : 2;
        s_par[0] = __int_as_float(a0); s_par[1] = __int_as_float(a1);
        s_par[2] = l[a0]; s_par[3] = l[a1];
        s_par[4] = 1.0f / fmaxf(ext[a0] / FS_G, 1e-30f); s_par[5] = 1.0f / fmaxf(ext[a1] / FS_G, 1e-30f);
    }
    __syncthreads();
    const int a0 = __float_as_int(s_par[0]), a1 = __float_as_int(s_par[1]);
    const float o0 = s_par[2], o1 = s_par[3], inv0 = s_par[4], inv1 = s_par[5];
    for (int k = tid; k < N; k += FS_T) {
        const float *p = pts + (size_t)k * stride;
        atomicAdd(&s_cnt[fs_cell(__ldg(p + a0), __ldg(p + a1), o0, o1, inv0, inv1)], 1);
    }
    __syncthreads();
    constexpr int PER = FS_CELLS / FS_T;
    int v[PER], sum = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) { v[i] = s_cnt[tid * PER + i]; sum += v[i]; }
    int incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = s_wsum[lane], wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, wi, d);
            if (lane >= d) wi += o;
        }
        s_wsum[lane] = wi - w;
    }
    __syncthreads();
    int run = s_wsum[warp] + incl - sum;
#pragma unroll
    for (int i = 0; i < PER; ++i) { s_cnt[tid * PER + i] = run; run += v[i]; }
    __syncthreads();
    for (int k = tid; k < N; k += FS_T) {
        const float *p = pts + (size_t)k * stride;
        const float x = __ldg(p), y = __ldg(p + 1), z = __ldg(p + 2);
        const float v0 = a0 == 0 ? x : (a0 == 1 ? y : z), v1 = a1 == 1 ? y : (a1 == 2 ? z : x);
        const int pos = atomicAdd(&s_cnt[fs_cell(v0, v1, o0, o1, inv0, inv1)], 1);
        sorted[pos] = make_float4(x, y, z, __int_as_float(k));
    }
}

#ifdef DCLR_TRACE
__device__ long long g_fps_trace[4096];
#define FTRACE(slot) do { if (blockIdx.x == 0 && lane == 0 && j >= 500 && j < 532) g_fps_trace[((j - 500) * 8 + warp) * 16 + (slot)] = clock64(); } while (0)
#define FTRACEV(slot, v) do { if (blockIdx.x == 0 && lane == 0 && j >= 500 && j < 532) g_fps_trace[((j - 500) * 8 + warp) * 16 + (slot)] = (v); } while (0)
extern "C" int dclr_debug_read_fps_trace(long long *dst, int n) {
    return cudaMemcpyFromSymbol(dst, g_fps_trace, sizeof(long long) * (n < 4096 ? n : 4096)) == cudaSuccess ? 0 : -1;
}
#else
#define FTRACE(slot) do { } while (0)
#define FTRACEV(slot, v) do { } while (0)
#endif

// NW warps, PP buckets per warp (bucket g = i * NW + warp); capacity CAP = 64 * NW * PP records.  Fewer, fatter warps
// win: per round every warp pays the skip test, two warp reductions and the barrier whether or not one of its buckets
// changes, while only ~9 of the 256 buckets of a 16384-point scan do (issue-bound at 32 warps, ncu r1b).
template <int NW, int PP>
__global__ void __launch_bounds__(NW * 32, 1)
fps_bucket_kernel(const float4 *__restrict__ sorted, int N, int npoint, int tie_ref, int32_t *__restrict__ idx_out,
                  float *__restrict__ new_xyz_out) {
    extern __shared__ __align__(16) float smem[];
    constexpr int CAP = 64 * NW * PP;
    constexpr int T = NW * 32;
    constexpr int WSTRIDE = 64 * NW;           // records between a warp's consecutive buckets
    static_assert(PP <= 32 && NW <= 32, "lane i keeps bucket i; lane w reads warp w's slot");
    float *s_x = smem, *s_y = smem + CAP, *s_z = smem + 2 * CAP;
    uint16_t *s_idx = reinterpret_cast<uint16_t *>(smem + 3 * CAP);             // tie rank of the original index
    int2 *s_slot = reinterpret_cast<int2 *>(s_idx + CAP);                       // [2][32] (max bits, key)
    int *s_start = reinterpret_cast<int *>(s_slot + 64);

    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    sorted += (size_t)b * N;
    idx_out += (size_t)b * npoint;
    if (new_xyz_out) new_xyz_out += (size_t)b * npoint * 3;

    for (int k = tid; k < CAP; k += T) {
        if (k < N) {
            const float4 r = __ldg(sorted + k);
            const int oi = __float_as_int(r.w);
            s_x[k] = r.x; s_y[k] = r.y; s_z[k] = r.z;
            s_idx[k] = (uint16_t)(tie_ref ? fb_rank16(oi) : oi);            // tie order (fb_rank16 above)
            if (oi == 0) *s_start = k;
        } else {
            s_x[k] = 0.f; s_y[k] = 0.f; s_z[k] = 0.f; s_idx[k] = 0xffffu;
        }
    }
    __syncthreads();

    // ---- per-thread running minima, per-bucket metadata (lane i < PP <-> the warp's i-th bucket)
    float tl[PP], th[PP];
    float blx = 0.f, bly = 0.f, blz = 0.f, bhx = 0.f, bhy = 0.f, bhz = 0.f, bm = -1.0f;
    uint32_t bkey = 0xffffffffu;
    const uint32_t rec0 = (uint32_t)(64 * warp + 2 * lane);                    // record of pair 0; pair i: + WSTRIDE * i
    const uint32_t ax = ptx::smem_u32(s_x) + rec0 * 4, ay = ptx::smem_u32(s_y) + rec0 * 4,
                   az = ptx::smem_u32(s_z) + rec0 * 4, ai = ptx::smem_u32(s_idx) + rec0 * 2;
#pragma unroll
    for (int i = 0; i < PP; ++i) {
        const int p = (int)rec0 + WSTRIDE * i;
        const bool v0 = p < N, v1 = p + 1 < N;
        tl[i] = v0 ? 1e10f : -1.0f;            // -1 can never win (reference: best starts at -1, strict >)
        th[i] = v1 ? 1e10f : -1.0f;
        float lo[3], hi[3];
        const float c0[3] = {s_x[p], s_y[p], s_z[p]}, c1[3] = {s_x[p + 1], s_y[p + 1], s_z[p + 1]};
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            lo[a] = fminf(v0 ? c0[a] : FLT_MAX, v1 ? c1[a] : FLT_MAX);
            hi[a] = fmaxf(v0 ? c0[a] : -FLT_MAX, v1 ? c1[a] : -FLT_MAX);
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) {
                lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], d));
                hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], d));
            }
        }
        const bool any = __any_sync(0xffffffffu, v0);
        if (lane == i) {
            blx = lo[0]; bly = lo[1]; blz = lo[2]; bhx = hi[0]; bhy = hi[1]; bhz = hi[2];
            bm = any ? 1e10f : -1.0f;          // 1e10 forces the first update; the key is set there
        }
    }

    int oldpos = *s_start;
    if (tid == 0) idx_out[0] = 0;
    for (int j = 1; j < npoint; ++j) {
        const float x1 = s_x[oldpos], y1 = s_y[oldpos], z1 = s_z[oldpos];
        if (tid == 0 && new_xyz_out) {
            new_xyz_out[(j - 1) * 3 + 0] = x1; new_xyz_out[(j - 1) * 3 + 1] = y1; new_xyz_out[(j - 1) * 3 + 2] = z1;
        }
        FTRACE(0);
        // ---- which of this warp's buckets can change?  (lane i tests bucket i)
        bool need = false;
        if (lane < PP) {
            const float ex = fmaxf(fmaxf(blx - x1, x1 - bhx), 0.f), ey = fmaxf(fmaxf(bly - y1, y1 - bhy), 0.f),
                        ez = fmaxf(fmaxf(blz - z1, z1 - bhz), 0.f);
            const float lb = (ex * ex + ey * ey + ez * ez) * 0.9999f;
            need = lb < bm;
        }
        // (vote.ballot beats redux.or here -- 0.534 vs 0.578 ms -- although only the latter yields a uniform register)
        const unsigned mask = __ballot_sync(0xffffffffu, need);
        FTRACE(1);
        FTRACEV(6, __popc(mask));
        if (mask) {
            const uint64_t xo = fb_pack2(x1, x1), yo = fb_pack2(y1, y1), zo = fb_pack2(z1, z1);
            // warp-uniform scan of the mask, a clean byte at a time; i is a compile-time constant so tl/th stay in
            // registers.  (Measured alternatives, tools/fps_trace.py + tools/micro/lat2.cu: a jump table on ffs(mask)
            // costs ~120 cycles per indirect branch and was slower, 0.71 vs 0.53 ms per 128 clouds.)
            auto upd = [&](const int i) {
                const uint64_t dx = fb_sub2(fb_lds64(ax + i * (WSTRIDE * 4)), xo),
                               dy = fb_sub2(fb_lds64(ay + i * (WSTRIDE * 4)), yo),
                               dz = fb_sub2(fb_lds64(az + i * (WSTRIDE * 4)), zo);
                const uint32_t id2 = fb_lds32(ai + i * (WSTRIDE * 2));
                const uint64_t d2 = fb_fma2(dz, dz, fb_fma2(dx, dx, fb_mul2(dy, dy)));
                float dl, dh;
                fb_unpack2(d2, dl, dh);
                tl[i] = fminf(dl, tl[i]);
                th[i] = fminf(dh, th[i]);
                const int bmx = __reduce_max_sync(0xffffffffu, __float_as_int(fmaxf(tl[i], th[i])));
                const uint32_t p = rec0 + (uint32_t)WSTRIDE * i;
                const uint32_t klo = __float_as_int(tl[i]) == bmx ? ((id2 & 0xffffu) << 16 | p) : 0xffffffffu;
                const uint32_t khi = __float_as_int(th[i]) == bmx ? ((id2 >> 16) << 16 | (p + 1)) : 0xffffffffu;
                const uint32_t kmin = __reduce_min_sync(0xffffffffu, min(klo, khi));
                if (lane == i) { bm = __int_as_float(bmx); bkey = kmin; }
            };
            // (a branch-free variant -- predicated selects over all PP register pairs -- was measured 2x slower)
            constexpr int GR = PP < 8 ? PP : 8;
#pragma unroll
            for (int q = 0; q < PP / GR; ++q) {
                if (PP > 8 && ((mask >> (q * GR)) & 0xffu) == 0) continue;
#pragma unroll
                for (int r = 0; r < GR; ++r)
                    if ((mask >> (q * GR + r)) & 1u) upd(q * GR + r);
            }
        }
        FTRACE(2);
        // ---- warp best over its buckets, block best over the warps (one barrier; slots double-buffered by parity)
        const int mybits = lane < PP ? __float_as_int(bm) : INT_MIN;
        const int wm = __reduce_max_sync(0xffffffffu, mybits);
        const uint32_t wk = __reduce_min_sync(0xffffffffu, (lane < PP && mybits == wm) ? bkey : 0xffffffffu);
        int2 *slot = s_slot + (j & 1) * 32;
        if (lane == 0) slot[warp] = make_int2(wm, (int)wk);
        FTRACE(3);
        __syncthreads();
        const int2 e = lane < NW ? slot[lane] : make_int2(INT_MIN, -1);
        const int m = __reduce_max_sync(0xffffffffu, e.x);
        const uint32_t key = __reduce_min_sync(0xffffffffu, e.x == m ? (uint32_t)e.y : 0xffffffffu);
        oldpos = (int)(key & 0xffffu);
        FTRACEV(4, clock64() + (oldpos & 0));
        if (tid == 0) idx_out[j] = tie_ref ? fb_unrank16(key >> 16) : (int)(key >> 16);
    }
    if (tid == 0 && new_xyz_out && npoint > 0) {
        new_xyz_out[(npoint - 1) * 3 + 0] = s_x[oldpos];
        new_xyz_out[(npoint - 1) * 3 + 1] = s_y[oldpos];
        new_xyz_out[(npoint - 1) * 3 + 2] = s_z[oldpos];
    }
}

// records [B*N] float4 (+ running minima [B*N] fp32 for the L2-resident variant, N > 16384)
size_t fps_bucket_workspace_bytes(int B, int N) {
    return (size_t)B * N * (sizeof(float4) + (N > 16384 ? sizeof(float) : 0)) + 256;
}

template <int NW, int PP>
static int launch_bucket(const float4 *sorted, int B, int N, int npoint, int tie_ref, int32_t *idx, float *nx,
                         cudaStream_t st) {
    const size_t smem = (size_t)64 * NW * PP * 14 + 64 * sizeof(int2) + 16;
    auto kern = fps_bucket_kernel<NW, PP>;
    DCLR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<B, NW * 32, smem, st>>>(sorted, N, npoint, tie_ref, idx, nx);
    DCLR_LAUNCH_CHECK("fps_bucket_kernel");
    return DCLR_OK;
}

// CAPK = capacity / 2048 records; warps x buckets per warp = CAPK * 32
template <int CAPK>
static int launch_bucket_any(int nw, const float4 *sorted, int B, int N, int npoint, int tie_ref, int32_t *idx, float *nx,
                             cudaStream_t st) {
    if (nw == 32) return launch_bucket<32, CAPK>(sorted, B, N, npoint, tie_ref, idx, nx, st);
    if (nw == 16) return launch_bucket<16, CAPK * 2>(sorted, B, N, npoint, tie_ref, idx, nx, st);
    return launch_bucket<8, CAPK * 4>(sorted, B, N, npoint, tie_ref, idx, nx, st);
}

// *used = 0 when the shape is not handled (caller falls back to fps_launch)
int fps_bucket_launch(const float *pts, int stride, int B, int N, int npoint, int tie_rule, int32_t *idx, float *nx,
                      void *ws, cudaStream_t st, int *used, cudaEvent_t after_sort) {
    *used = 0;
    const int tie_ref = tie_rule == DCLR_TIE_REFERENCE;
    const char *env_min = getenv("DCLR_FPS_BUCKET_MIN");          // tuning knob (read per call: tests switch it); measured
    const int min_n = env_min ? atoi(env_min) : 8192;
    // per 8 clouds, bucketed | resident kernel: N = 4096 0.154 | 0.118 ms, 8192 0.306 | 0.295, 12288 0.416 | 0.727
    if (N <= min_n || N > 131072 || npoint < 2) return DCLR_OK;
    float4 *sorted = (float4 *)ws;
    fps_sort_kernel<<<B, FS_T, 0, st>>>(pts, stride, N, sorted);
    DCLR_LAUNCH_CHECK("fps_sort_kernel");
    if (after_sort) DCLR_CUDA(cudaEventRecord(after_sort, st));   // the sampling kernel leaves SMs idle: see model.cu
    int rc;
    if (N > 16384) {                    // point state in L2 (fps_bucket_l2.cu); block of the reference kernel = 1024
        rc = fps_bucket_l2_launch(sorted, (float *)(sorted + (size_t)B * N), B, N, npoint, tie_ref ? 10 : -1, idx, nx, st);
        if (rc == DCLR_OK) *used = 1;
        return rc;
    }
    // tuning knobs, read per call (tools/fps_ab.py, tests): DCLR_FPS_SHAPE = warps per cloud; DCLR_FPS_MP = samples per
    // block-wide round at most; DCLR_FPS_MP=0 = the register-resident single-pick kernel of this file
    const char *env_shape = getenv("DCLR_FPS_SHAPE"), *env_mp = getenv("DCLR_FPS_MP");
    const int shape = env_shape ? atoi(env_shape) : 8;
    const int mp = env_mp ? atoi(env_mp) : 3;      // measured per 128 clouds: 1 0.567 | 2 0.457 | 3 0.428 | 4 0.435 | 6 0.490 ms
    const char *env_cl = getenv("DCLR_FPS_CLUSTER");        // 1: one cloud over a cluster of two CTAs (fps_cluster.cu)
    const int use_cluster = env_cl ? atoi(env_cl) : 0;
    if (mp >= 1 && use_cluster && N > 8192) rc = fps_cluster_launch(sorted, B, N, npoint, tie_ref, mp, idx, nx, st);
    else if (mp >= 1) rc = fps_tmem_launch(sorted, B, N, npoint, tie_ref, shape, mp, idx, nx, st);
    else if (N <= 4096) rc = launch_bucket_any<2>(shape, sorted, B, N, npoint, tie_ref, idx, nx, st);
    else if (N <= 8192) rc = launch_bucket_any<4>(shape, sorted, B, N, npoint, tie_ref, idx, nx, st);
    else rc = launch_bucket_any<8>(shape, sorted, B, N, npoint, tie_ref, idx, nx, st);
    if (rc == DCLR_OK) *used = 1;
    return rc;
}

}  // namespace dclr
