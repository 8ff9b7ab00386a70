// Bucketed farthest point sampling for clouds that do not fit one SM's shared memory (16384 < N <= 131072): the same
// exact pruning as fps_bucket.cu (Z-ordered records, 64-record buckets, "box lower bound >= bucket max t => skip"), but
// the per-point state -- records (x, y, z, index) and the running minima t -- stays in global memory, i.e. in L2:
// a round touches a handful of buckets (~1 KB each), not the cloud, so the L2 round trip is paid per touched bucket
// instead of the N x 16 B per round that the streaming kernel (fps.cu) and the reference kernel
// (sampling_gpu.cu:113-208) move.  Per-bucket metadata (box, max t, tie rank / position / coordinates of that maximum)
// lives in registers: thread <-> bucket, up to two buckets per thread.  Arithmetic and tie rules as everywhere
// (fma(dz,dz,fma(dx,dx,dy*dy)); lowest index or the reference's tree order), so the indices are bit-identical to the
// other FPS kernels and to the reference.
#include "common.cuh"
#include "kernels.cuh"
#include <float.h>
#include <limits.h>

namespace dclr {

constexpr int FL_T = 1024;                 // threads per CTA (one CTA per cloud)
constexpr int FL_BS = 64;                  // records per bucket: lane l owns records 2l, 2l+1

__device__ __forceinline__ int fl_rank(int k, int L) {             // same order as fps.cu:tie_rank
    if (L <= 0) return k;
    return (int)(((__brev((uint32_t)k & ((1u << L) - 1u)) >> (32 - L)) << (31 - L)) | ((uint32_t)k >> L));
}
__device__ __forceinline__ int fl_unrank(int r, int L) {
    if (L <= 0) return r;
    const uint32_t slot = __brev((uint32_t)r >> (31 - L)) >> (32 - L), q = (uint32_t)r & ((1u << (31 - L)) - 1u);
    return (int)((q << L) | slot);
}

struct FlSlot { int m; int rank; float x, y, z; int pad[3]; };     // 32 bytes

// PB buckets per thread.  Bucket g = s * 1024 + lane * 32 + warp  (s < PB): consecutive buckets sit on different warps.
template <int PB>
__global__ void __launch_bounds__(FL_T, 1)
fps_bucket_l2_kernel(const float4 *__restrict__ sorted, float *__restrict__ tmin, int N, int npoint, int tieL,
                     int32_t *__restrict__ idx_out, float *__restrict__ new_xyz_out) {
    __shared__ FlSlot s_slot[2][32];
    __shared__ float s_first[4];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    sorted += (size_t)b * N;
    tmin += (size_t)b * N;
    idx_out += (size_t)b * npoint;
    if (new_xyz_out) new_xyz_out += (size_t)b * npoint * 3;
    const int nbk = (N + FL_BS - 1) / FL_BS;

    // ---- metadata of my buckets (lane i of this warp <-> bucket s*1024 + i*32 + warp); t starts at 1e10
    float blx[PB], bly[PB], blz[PB], bhx[PB], bhy[PB], bhz[PB], bm[PB], bx[PB], by[PB], bz[PB];
    int brank[PB];
#pragma unroll
    for (int s = 0; s < PB; ++s) { bm[s] = -1.0f; brank[s] = INT_MAX; bx[s] = by[s] = bz[s] = 0.f;
                                   blx[s] = bly[s] = blz[s] = bhx[s] = bhy[s] = bhz[s] = 0.f; }
#pragma unroll
    for (int s = 0; s < PB; ++s)
        for (int i = 0; i < 32; ++i) {
            const int g = s * 1024 + i * 32 + warp;
            if (g >= nbk) break;                                    // warp-uniform
            const int p = g * FL_BS + 2 * lane;
            float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
#pragma unroll
            for (int e = 0; e < 2; ++e)
                if (p + e < N) {
                    const float4 r = __ldg(sorted + p + e);
                    lo[0] = fminf(lo[0], r.x); hi[0] = fmaxf(hi[0], r.x);
                    lo[1] = fminf(lo[1], r.y); hi[1] = fmaxf(hi[1], r.y);
                    lo[2] = fminf(lo[2], r.z); hi[2] = fmaxf(hi[2], r.z);
                    tmin[p + e] = 1e10f;
                    if (__float_as_int(r.w) == 0) { s_first[0] = r.x; s_first[1] = r.y; s_first[2] = r.z; }
                }
#pragma unroll
            for (int a = 0; a < 3; ++a)
#pragma unroll
                for (int d = 16; d >= 1; d >>= 1) {
                    lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], d));
                    hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], d));
                }
            if (lane == i) {
                blx[s] = lo[0]; bly[s] = lo[1]; blz[s] = lo[2]; bhx[s] = hi[0]; bhy[s] = hi[1]; bhz[s] = hi[2];
                bm[s] = 1e10f;                                      // forces the first update
            }
        }
    __syncthreads();

    float x1 = s_first[0], y1 = s_first[1], z1 = s_first[2];
    if (tid == 0) idx_out[0] = 0;
    for (int j = 1; j < npoint; ++j) {
        if (tid == 0 && new_xyz_out) {
            new_xyz_out[(j - 1) * 3 + 0] = x1; new_xyz_out[(j - 1) * 3 + 1] = y1; new_xyz_out[(j - 1) * 3 + 2] = z1;
        }
#pragma unroll
        for (int s = 0; s < PB; ++s) {
            const float ex = fmaxf(fmaxf(blx[s] - x1, x1 - bhx[s]), 0.f), ey = fmaxf(fmaxf(bly[s] - y1, y1 - bhy[s]), 0.f),
                        ez = fmaxf(fmaxf(blz[s] - z1, z1 - bhz[s]), 0.f);
            const bool need = (ex * ex + ey * ey + ez * ez) * 0.9999f < bm[s];
            unsigned rest = __ballot_sync(0xffffffffu, need);
            while (rest) {                                          // warp-uniform: one touched bucket at a time
                const int i = __ffs((int)rest) - 1;
                rest &= rest - 1;
                const int p = (s * 1024 + i * 32 + warp) * FL_BS + 2 * lane;
                float t0 = -1.0f, t1 = -1.0f;
                int r0 = INT_MAX, r1 = INT_MAX;
                float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f), q1 = q0;
                if (p + 1 < N) {
                    q0 = __ldcg(sorted + p); q1 = __ldcg(sorted + p + 1);
                    const float ta = __ldcg(tmin + p), tb = __ldcg(tmin + p + 1);
                    t0 = fminf(sqdist_pn2(q0.x - x1, q0.y - y1, q0.z - z1), ta);
                    t1 = fminf(sqdist_pn2(q1.x - x1, q1.y - y1, q1.z - z1), tb);
                    if (t0 != ta) __stcg(tmin + p, t0);
                    if (t1 != tb) __stcg(tmin + p + 1, t1);
                } else if (p < N) {
                    q0 = __ldcg(sorted + p);
                    const float tt = __ldcg(tmin + p);
                    t0 = fminf(sqdist_pn2(q0.x - x1, q0.y - y1, q0.z - z1), tt);
                    if (t0 != tt) __stcg(tmin + p, t0);
                }
                const int bmx = __reduce_max_sync(0xffffffffu, __float_as_int(fmaxf(t0, t1)));
                if (__float_as_int(t0) == bmx) r0 = fl_rank(__float_as_int(q0.w), tieL);
                if (__float_as_int(t1) == bmx) r1 = fl_rank(__float_as_int(q1.w), tieL);
                const int rmin = (int)__reduce_min_sync(0xffffffffu, (unsigned)min(r0, r1));
                // the unique lane holding rank rmin hands the coordinates to the bucket's owner lane i
                const bool mine0 = r0 == rmin, mine1 = r1 == rmin;
                const unsigned hold = __ballot_sync(0xffffffffu, (mine0 || mine1) && rmin != INT_MAX);
                const int src = hold ? __ffs((int)hold) - 1 : 0;
                const float hx = __shfl_sync(0xffffffffu, mine0 ? q0.x : q1.x, src);
                const float hy = __shfl_sync(0xffffffffu, mine0 ? q0.y : q1.y, src);
                const float hz = __shfl_sync(0xffffffffu, mine0 ? q0.z : q1.z, src);
                if (lane == i) { bm[s] = __int_as_float(bmx); brank[s] = rmin; bx[s] = hx; by[s] = hy; bz[s] = hz; }
            }
        }
        // ---- warp best over its buckets -> slot; block best over the warps (one barrier, slots double-buffered)
        int mb = __float_as_int(bm[0]), mr = brank[0];
        float mx = bx[0], my = by[0], mz = bz[0];
#pragma unroll
        for (int s = 1; s < PB; ++s) {
            const int c = __float_as_int(bm[s]);
            if (c > mb || (c == mb && brank[s] < mr)) { mb = c; mr = brank[s]; mx = bx[s]; my = by[s]; mz = bz[s]; }
        }
        const int wm = __reduce_max_sync(0xffffffffu, mb);
        const int wr = (int)__reduce_min_sync(0xffffffffu, mb == wm ? (unsigned)mr : 0xffffffffu);
        FlSlot *slot = s_slot[j & 1];
        if (mb == wm && mr == wr) { slot[warp].m = wm; slot[warp].rank = wr; slot[warp].x = mx; slot[warp].y = my; slot[warp].z = mz; }
        __syncthreads();
        const int em = slot[lane].m, er = slot[lane].rank;
        const int m = __reduce_max_sync(0xffffffffu, em);
        const int r = (int)__reduce_min_sync(0xffffffffu, em == m ? (unsigned)er : 0xffffffffu);
        const unsigned wh = __ballot_sync(0xffffffffu, em == m && er == r);
        const int wsrc = __ffs((int)wh) - 1;
        x1 = slot[wsrc].x; y1 = slot[wsrc].y; z1 = slot[wsrc].z;
        if (tid == 0) idx_out[j] = fl_unrank(r, tieL);
    }
    if (tid == 0 && new_xyz_out && npoint > 0) {
        new_xyz_out[(npoint - 1) * 3 + 0] = x1; new_xyz_out[(npoint - 1) * 3 + 1] = y1; new_xyz_out[(npoint - 1) * 3 + 2] = z1;
    }
}

int fps_bucket_l2_launch(const float4 *sorted, float *tmin, int B, int N, int npoint, int tieL, int32_t *idx, float *nx,
                         cudaStream_t st) {
    if (N <= 65536) fps_bucket_l2_kernel<1><<<B, FL_T, 0, st>>>(sorted, tmin, N, npoint, tieL, idx, nx);
    else fps_bucket_l2_kernel<2><<<B, FL_T, 0, st>>>(sorted, tmin, N, npoint, tieL, idx, nx);
    DCLR_LAUNCH_CHECK("fps_bucket_l2_kernel");
    return DCLR_OK;
}

}  // namespace dclr
