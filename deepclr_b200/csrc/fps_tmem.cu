// Bucketed farthest point sampling with the running minima in TENSOR MEMORY and several samples per block-wide round.
// Same exact pruning as fps_bucket.cu (Z-ordered cloud, 64-point buckets, box lower bound), same arithmetic and tie
// order (sampling_gpu.cu:113-208 of the reference), two changes to what a round costs:
//
// 1. The running minima t_k (64 per thread) live in TMEM instead of registers.  Registers cannot be indexed at run
//    time, so fps_bucket.cu carries one unrolled copy of the update per bucket (32 per warp, ~20 KB of code) and walks a
//    chain of branches to the touched ones: i-cache misses and taken branches cost more than the update itself (~350 of
//    ~960 cycles per round).  TMEM is 256 KB per SM of storage that is private to a lane AND addressed by a RUNTIME column
//    (tcgen05.ld / tcgen05.st .32x32b: lane l of warp w reads its own row l of lane quarter w % 4).  With the state of
//    bucket i in columns 4i .. 4i+2 (t of the lane's two points and their two 16-bit tie ranks: one tcgen05.ld.x4) the
//    update is ONE generic loop body over the set bits of the skip mask, two buckets per iteration so that their
//    load -> distance -> redux chains overlap; no tensor-core instruction is involved.  Shared memory holds only the
//    coordinates (193 KB), which leaves ~30 KB of the SM to small kernels of other streams.
//
// 2. Up to MP samples per round, EXACTLY.  At the barrier every warp w publishes its best point (value v_w, tie key) and
//    an upper bound u_w of every OTHER running minimum in the warp.  Sort the NW warp winners c_0 >= c_1 >= ... by
//    (value, tie key).  c_0 is the next sample.  c_r (r >= 1) is provably the sample after c_0 .. c_{r-1} when for every
//    l < r:  d(c_r, c_l) >= t[c_r]  (its minimum is untouched by the picks before it: min(d, t) == t),
//            u_{warp(c_l)} < t[c_r] (nothing that is left in an earlier winner's warp can overtake it) and t[c_r] > 0.
//    Every other minimum can only fall; the points of warps ranked after c_r were <= t[c_r] with a later tie key
//    already, the rest of c_r's own warp likewise, the rest of the earlier winners' warps is < t[c_r] by u, the earlier
//    picks themselves drop to 0.  The accepted prefix is applied in ONE update pass (t = min(t, d_0, .., d_{MP-1});
//    min-updates commute; unused pick slots sit at 1e18 and change nothing).  tools/fps_multipick_sim.py: the 1023
//    serial rounds of a KITTI-shaped scan become ~565 (MP = 2), ~427 (3), ~380 (4).  Per bucket the kernel keeps
//    bm2 = an upper bound of its second-largest minimum for u (one more redux per update).
#include "common.cuh"
#include "kernels.cuh"
#include "ptx.cuh"
#include "fps_bucket.cuh"
#include <float.h>
#include <limits.h>

namespace dclr {

#ifdef DCLR_TRACE
__device__ long long g_fps_trace_tm[4096];
#define MTRACE(slot) do { if (blockIdx.x == 0 && lane == 0 && round >= 100 && round < 132) g_fps_trace_tm[((round - 100) * 8 + (warp & 7)) * 16 + (slot)] = clock64(); } while (0)
#define MTRACEV(slot, v) do { if (blockIdx.x == 0 && lane == 0 && round >= 100 && round < 132) g_fps_trace_tm[((round - 100) * 8 + (warp & 7)) * 16 + (slot)] = (v); } while (0)
extern "C" int dclr_debug_read_fps_trace_tm(long long *dst, int n) {
    return cudaMemcpyFromSymbol(dst, g_fps_trace_tm, sizeof(long long) * (n < 4096 ? n : 4096)) == cudaSuccess ? 0 : -1;
}
#else
#define MTRACE(slot) do { } while (0)
#define MTRACEV(slot, v) do { } while (0)
#endif

__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}

// NW warps (multiple of 4: warps w, w + 4, .. share TMEM lane quarter w % 4), PP buckets per warp (bucket g = i * NW +
// warp, lane i keeps its metadata), MP samples per round at most.
template <int NW, int PP, int MP>
__global__ void __launch_bounds__(NW * 32, 1)
fps_tmem_kernel(const float4 *__restrict__ sorted, int N, int npoint, int tie_ref, int32_t *__restrict__ idx_out,
                float *__restrict__ new_xyz_out) {
    extern __shared__ __align__(16) float smem[];
    constexpr int CAP = 64 * NW * PP;
    constexpr int WSTRIDE = 64 * NW;
    constexpr int NPAIR = MP * (MP - 1) / 2;
    constexpr int TCOLS = (NW / 4) * 4 * PP < 32 ? 32 : (NW / 4) * 4 * PP;     // 256 for 16384 records: {t lo, t hi, tie ranks, -}
    static_assert(NW % 4 == 0 && PP <= 32 && NW <= 32 && MP >= 1 && MP <= NW && NPAIR <= 32, "shape");
    static_assert((TCOLS & (TCOLS - 1)) == 0 && TCOLS <= 512, "TMEM columns: power of two");
    float *s_x = smem, *s_y = smem + CAP, *s_z = smem + 2 * CAP;
    // warp entries [2][NW] x {value^sign, ~key, bound, -}, {x, y, z, -}; per warp the first MP winners in order [NW][MP] x same
    int4 *s_slot = reinterpret_cast<int4 *>(smem + 3 * CAP);
    int4 *s_cand = s_slot + 4 * NW;
    int *s_start = reinterpret_cast<int *>(s_cand);                              // (both only live before the first round)
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(s_cand) + 1;

    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    sorted += (size_t)b * N;
    idx_out += (size_t)b * npoint;
    if (new_xyz_out) new_xyz_out += (size_t)b * npoint * 3;

    if (warp == 0) ptx::tmem_alloc(s_tmem, TCOLS);
    if (tid == 0) *s_start = 0;
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    // this thread's state of bucket i: TMEM lane (32 (warp % 4) + lane), columns tcol + 4 i: t of its two points and their
    // tie ranks.  (The ranks used to sit in 32 KB of shared memory; without them the CTA leaves ~30 KB of the SM to the
    // small kernels of other streams, which can then run beside this latency-bound one.)
    const uint32_t tmem_base = *s_tmem;
    const uint32_t tcol = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * 4 * PP);

    // ---- stage the cloud (every lane loads the two records it owns: 1 KB contiguous per warp and bucket, 8 loads in flight)
    //      + per-bucket metadata (lane i < PP <-> the warp's i-th bucket): box, max t, key of the max, second max (bound)
    float blx = 0.f, bly = 0.f, blz = 0.f, bhx = 0.f, bhy = 0.f, bhz = 0.f, bm = -1.0f, bm2 = -1.0f;
    uint32_t bkey = 0xffffff00u | (uint32_t)warp;
    const uint32_t rec0 = (uint32_t)(64 * warp + 2 * lane);                    // record of pair 0; pair i: + WSTRIDE * i
    const uint32_t ax = ptx::smem_u32(s_x) + rec0 * 4, ay = ptx::smem_u32(s_y) + rec0 * 4,
                   az = ptx::smem_u32(s_z) + rec0 * 4;
    constexpr int UNR = PP % 4 == 0 ? 4 : 1;
    for (int i0 = 0; i0 < PP; i0 += UNR) {
        float4 ra[UNR], rb[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const int p = (int)rec0 + WSTRIDE * (i0 + u);
            ra[u] = p < N ? __ldg(sorted + p) : make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
            rb[u] = p + 1 < N ? __ldg(sorted + p + 1) : make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const int i = i0 + u;
            const int p = (int)rec0 + WSTRIDE * i;
            const int o0 = __float_as_int(ra[u].w), o1 = __float_as_int(rb[u].w);
            const bool v0 = o0 >= 0, v1 = o1 >= 0;
            s_x[p] = ra[u].x; s_y[p] = ra[u].y; s_z[p] = ra[u].z;
            s_x[p + 1] = rb[u].x; s_y[p + 1] = rb[u].y; s_z[p + 1] = rb[u].z;
            if (o0 == 0) *s_start = p;
            if (o1 == 0) *s_start = p + 1;
            const uint32_t r0 = v0 ? (uint32_t)(tie_ref ? fb_rank16(o0) : o0) & 0xffffu : 0xffffu;     // tie order (fb_rank16)
            const uint32_t r1 = v1 ? (uint32_t)(tie_ref ? fb_rank16(o1) : o1) & 0xffffu : 0xffffu;
            // -1 can never win (reference: best starts at -1, strict >)
            ptx::tmem_st4(tcol + 4 * i, __float_as_uint(v0 ? 1e10f : -1.0f), __float_as_uint(v1 ? 1e10f : -1.0f),
                          r0 | (r1 << 16), 0u);
            float lo[3], hi[3];
            const float c0[3] = {ra[u].x, ra[u].y, ra[u].z}, c1[3] = {rb[u].x, rb[u].y, rb[u].z};
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
                bm = any ? 1e10f : -1.0f;          // 1e10 forces the first update; key and bm2 are set there
                bm2 = bm;
            }
        }
    }
    __syncthreads();                               // *s_start and everybody's coordinates
    ptx::tmem_st_wait();

    // candidate pair (pl, pr), pl < pr, tested by this lane: pair index p = pr (pr - 1) / 2 + pl
    int pl = 0, pr = 1;
#pragma unroll
    for (int r = 1; r < MP; ++r)
        if (lane >= r * (r - 1) / 2) { pr = r; pl = lane - r * (r - 1) / 2; }
    const bool pair_lane = lane < NPAIR;
    if (!pair_lane) { pl = 0; pr = MP > 1 ? 1 : 0; }       // (idle lanes read a valid pair; their vote is masked)
    const int4 *cand = s_cand + warp * (2 * MP);  // this warp's copy of the first MP winners in order: {hi, lo, bound, -}, {x, y, z, -}

    // picks applied in the coming round: coordinates in registers (unused slots far away: they change nothing)
    float px[MP], py[MP], pz[MP];
#pragma unroll
    for (int q = 0; q < MP; ++q) { px[q] = 1e18f; py[q] = 1e18f; pz[q] = 1e18f; }
    {
        const int st = *s_start;
        px[0] = s_x[st]; py[0] = s_y[st]; pz[0] = s_z[st];
        if (tid == 0) {
            idx_out[0] = 0;
            if (new_xyz_out) { new_xyz_out[0] = px[0]; new_xyz_out[1] = py[0]; new_xyz_out[2] = pz[0]; }
        }
    }
    int done = 1;                                 // samples selected (and written) so far
    uint32_t ehi = 0u, elo = 0u;                  // this warp's published entry (recomputed only when a bucket changed)
    int ew2 = INT_MIN;
    float ex = 0.f, ey = 0.f, ez = 0.f;
    bool efresh = false;
    for (int round = 0; done < npoint; ++round) {
        MTRACE(0);
        // ---- which of this warp's buckets can change?  (lane i tests bucket i against every pick; branch-free)
        unsigned need = 0u;
#pragma unroll
        for (int q = 0; q < MP; ++q) {
            const float dx = fmax3(blx - px[q], px[q] - bhx, 0.f), dy = fmax3(bly - py[q], py[q] - bhy, 0.f),
                        dz = fmax3(blz - pz[q], pz[q] - bhz, 0.f);
            need |= (unsigned)((dx * dx + dy * dy + dz * dz) * 0.9999f < bm);
        }
        unsigned mask = __ballot_sync(0xffffffffu, need != 0u);     // (lanes >= PP keep bm = -1: never set)
        MTRACE(1);
        MTRACEV(6, __popc(mask));
        const bool touched = mask != 0u;
        if (touched) {
            uint64_t xo[MP], yo[MP], zo[MP];
#pragma unroll
            for (int q = 0; q < MP; ++q) { xo[q] = fb_pack2(px[q], px[q]); yo[q] = fb_pack2(py[q], py[q]); zo[q] = fb_pack2(pz[q], pz[q]); }
            ptx::tmem_st_wait();                  // the stores of the previous round
            while (mask) {
                // two buckets per iteration (the second repeats the first when only one is left: same values again)
                int iu[2];
                iu[0] = 31 - __clz(mask);               // highest first: one FLO (ffs = BREV + FLO, ~47 dependent cycles)
                mask ^= 1u << iu[0];
                iu[1] = mask ? 31 - __clz(mask) : iu[0];
                mask &= ~(1u << iu[1]);
                uint64_t X[2], Y[2], Z[2];
                uint32_t id2[2], tlb[2], thb[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    uint32_t pad;
                    X[u] = fb_lds64(ax + iu[u] * (WSTRIDE * 4)); Y[u] = fb_lds64(ay + iu[u] * (WSTRIDE * 4));
                    Z[u] = fb_lds64(az + iu[u] * (WSTRIDE * 4));
                    ptx::tmem_ld4(tcol + 4 * iu[u], tlb[u], thb[u], id2[u], pad);
                }
                ptx::tmem_ld_wait();
                int mx[2], mn[2], nlb[2], nhb[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    float nl = __uint_as_float(tlb[u]), nh = __uint_as_float(thb[u]);
#pragma unroll
                    for (int q = 0; q < MP; ++q) {
                        // the reference's arithmetic (sampling_gpu.cu:133-137 as contracted), two points per instruction
                        const uint64_t dx = fb_sub2(X[u], xo[q]), dy = fb_sub2(Y[u], yo[q]), dz = fb_sub2(Z[u], zo[q]);
                        const uint64_t d2 = fb_fma2(dz, dz, fb_fma2(dx, dx, fb_mul2(dy, dy)));
                        float dl, dh;
                        fb_unpack2(d2, dl, dh);
                        nl = fminf(dl, nl);
                        nh = fminf(dh, nh);
                    }
                    ptx::tmem_st2(tcol + 4 * iu[u], __float_as_uint(nl), __float_as_uint(nh));
                    nlb[u] = __float_as_int(nl); nhb[u] = __float_as_int(nh);
                    mx[u] = max(nlb[u], nhb[u]);          // integer images: values are >= 0 or exactly -1.0f (negative int)
                    mn[u] = min(nlb[u], nhb[u]);
                }
                int bmx[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) bmx[u] = __reduce_max_sync(0xffffffffu, mx[u]);
                uint32_t kmin[2];
                int b2[2];
                unsigned hb[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const uint32_t p = rec0 + (uint32_t)WSTRIDE * iu[u];
                    const uint32_t klo = nlb[u] == bmx[u] ? ((id2[u] & 0xffffu) << 16 | p) : 0xffffffffu;
                    const uint32_t khi = nhb[u] == bmx[u] ? ((id2[u] >> 16) << 16 | (p + 1)) : 0xffffffffu;
                    const bool holder = mx[u] == bmx[u];
                    kmin[u] = __reduce_min_sync(0xffffffffu, min(klo, khi));
                    // second-largest minimum of the bucket (upper bound): the holder lane contributes its other point;
                    // two holder lanes (or both points of one lane) => the maximum itself
                    b2[u] = __reduce_max_sync(0xffffffffu, holder ? mn[u] : mx[u]);
                    hb[u] = __ballot_sync(0xffffffffu, holder);
                }
#pragma unroll
                for (int u = 0; u < 2; ++u)
                    if (lane == iu[u]) {
                        bm = __int_as_float(bmx[u]);
                        bkey = bmx[u] >= 0 ? kmin[u] : (0xffffff00u | (uint32_t)warp);
                        bm2 = __int_as_float(__popc(hb[u]) > 1 ? bmx[u] : b2[u]);
                    }
            }
        }
        MTRACE(2);
        // ---- this warp's entry: best point (value, key, coordinates), bound of everything else in the warp
        if (touched || !efresh) {
            const int mybits = lane < PP ? __float_as_int(bm) : INT_MIN;
            const int ewm = __reduce_max_sync(0xffffffffu, mybits);
            const bool iswin = lane < PP && mybits == ewm;
            const uint32_t ewk = __reduce_min_sync(0xffffffffu, iswin ? bkey : 0xffffffffu);
            ew2 = __reduce_max_sync(0xffffffffu, iswin ? __float_as_int(bm2) : mybits);
            if (__popc(__ballot_sync(0xffffffffu, iswin)) > 1) ew2 = ewm;
            ehi = (uint32_t)ewm ^ 0x80000000u;            // larger = better, as one unsigned 64-bit number (ehi:elo)
            elo = ~ewk;
            const int wp = ewm >= 0 ? (int)(ewk & 0xffffu) : 0;     // (an invalid winner's key holds no position)
            ex = s_x[wp]; ey = s_y[wp]; ez = s_z[wp];
            efresh = true;
        }
        int4 *slot = s_slot + (round & 1) * (2 * NW);
        if (lane == 0) {
            slot[2 * warp] = make_int4((int)ehi, (int)elo, ew2, 0);
            slot[2 * warp + 1] = make_int4(__float_as_int(ex), __float_as_int(ey), __float_as_int(ez), 0);
        }
        MTRACE(3);
        __syncthreads();
        // ---- every warp ranks the NW entries (value desc, key asc; keys are distinct) and lays the first MP out in order
        {
            const int me_i = lane < NW ? lane : 0;
            const int4 eA = slot[2 * me_i], eB = slot[2 * me_i + 1];
            const uint64_t me = ((uint64_t)(uint32_t)eA.x << 32) | (uint32_t)eA.y;
            int c[NW];
#pragma unroll
            for (int a = 0; a < NW; ++a) {
                const uint2 o = *reinterpret_cast<const uint2 *>(slot + 2 * a);
                c[a] = (int)((((uint64_t)o.x << 32) | o.y) > me);
            }
#pragma unroll
            for (int s = 1; s < NW; s <<= 1)
#pragma unroll
                for (int a = 0; a + s < NW; a += 2 * s) c[a] += c[a + s];
            if (lane < NW && c[0] < MP) {
                s_cand[warp * (2 * MP) + 2 * c[0]] = eA;
                s_cand[warp * (2 * MP) + 2 * c[0] + 1] = eB;
            }
        }
        __syncwarp();
        // the MP winners' coordinates, loaded before it is known how many are taken (off the vote's critical path)
        uint32_t cx[MP], cy[MP], cz[MP];
        const uint32_t cand_a = ptx::smem_u32(cand);
#pragma unroll
        for (int q = 0; q < MP; ++q) {
            uint32_t pad;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(cx[q]), "=r"(cy[q]), "=r"(cz[q]), "=r"(pad)
                         : "r"(cand_a + (2 * q + 1) * 16));
        }
        int take = 1;
        if (MP > 1) {
            // lane p tests pair (pl, pr): the update's own arithmetic, point c_pr against sample c_pl
            const int4 lA = cand[2 * pl], lB = cand[2 * pl + 1], rA = cand[2 * pr], rB = cand[2 * pr + 1];
            const int rv = (int)((uint32_t)rA.x ^ 0x80000000u);
            const float dx = __int_as_float(rB.x) - __int_as_float(lB.x), dy = __int_as_float(rB.y) - __int_as_float(lB.y),
                        dz = __int_as_float(rB.z) - __int_as_float(lB.z);
            const float d = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
            const bool ok = pair_lane & (rv > 0) & (d >= __int_as_float(rv)) & (lA.z < rv);
            const unsigned okm = __ballot_sync(0xffffffffu, ok);
#pragma unroll
            for (int r = 1; r < MP; ++r) {
                const unsigned want = ((1u << r) - 1u) << (r * (r - 1) / 2);
                take += (int)((take == r) & ((okm & want) == want));
            }
            take = min(take, npoint - done);
        }
        // next round's picks (uniform) ...
#pragma unroll
        for (int q = 0; q < MP; ++q) {
            const bool on = q < take;
            px[q] = on ? __uint_as_float(cx[q]) : 1e18f; py[q] = on ? __uint_as_float(cy[q]) : 1e18f;
            pz[q] = on ? __uint_as_float(cz[q]) : 1e18f;
        }
        MTRACEV(4, clock64() + (take & 0));
        MTRACEV(7, take);
        // ... and their record: lane q of warp 0 writes sample done + q
        if (tid < take) {
            const int4 cA = cand[2 * tid], cB = cand[2 * tid + 1];
            const uint32_t key = ~(uint32_t)cA.y;
            idx_out[done + tid] = tie_ref ? fb_unrank16(key >> 16) : (int)(key >> 16);
            if (new_xyz_out) {
                float *o = new_xyz_out + (size_t)(done + tid) * 3;
                o[0] = __int_as_float(cB.x); o[1] = __int_as_float(cB.y); o[2] = __int_as_float(cB.z);
            }
        }
        done += take;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc(tmem_base, TCOLS);
}

template <int NW, int PP, int MP>
static int launch_tmem(const float4 *sorted, int B, int N, int npoint, int tie_ref, int32_t *idx, float *nx,
                       cudaStream_t st) {
    const size_t smem = (size_t)64 * NW * PP * 12 + (size_t)(4 * NW + 2 * NW * MP) * sizeof(int4);
    static_assert((size_t)64 * NW * PP * 12 + (size_t)(4 * NW + 2 * NW * MP) * sizeof(int4) <= 232448, "shared memory");
    auto kern = fps_tmem_kernel<NW, PP, MP>;
    DCLR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<B, NW * 32, smem, st>>>(sorted, N, npoint, tie_ref, idx, nx);
    DCLR_LAUNCH_CHECK("fps_tmem_kernel");
    return DCLR_OK;
}

// CAPK = capacity / 2048 records (warps x buckets per warp = CAPK * 32); nw warps per cloud; mp samples per round
template <int CAPK>
static int launch_tmem_any(int nw, int mp, const float4 *sorted, int B, int N, int npoint, int tie_ref, int32_t *idx,
                           float *nx, cudaStream_t st) {
    if (nw == 16) {
        if (mp <= 1) return launch_tmem<16, CAPK * 2, 1>(sorted, B, N, npoint, tie_ref, idx, nx, st);
        if (mp == 2) return launch_tmem<16, CAPK * 2, 2>(sorted, B, N, npoint, tie_ref, idx, nx, st);
        return launch_tmem<16, CAPK * 2, 4>(sorted, B, N, npoint, tie_ref, idx, nx, st);      // (shared memory: 16 x MP entries)
    }
    if (mp <= 1) return launch_tmem<8, CAPK * 4, 1>(sorted, B, N, npoint, tie_ref, idx, nx, st);
    if (mp == 2) return launch_tmem<8, CAPK * 4, 2>(sorted, B, N, npoint, tie_ref, idx, nx, st);
    if (mp == 3) return launch_tmem<8, CAPK * 4, 3>(sorted, B, N, npoint, tie_ref, idx, nx, st);
    if (mp <= 4) return launch_tmem<8, CAPK * 4, 4>(sorted, B, N, npoint, tie_ref, idx, nx, st);
    return launch_tmem<8, CAPK * 4, 6>(sorted, B, N, npoint, tie_ref, idx, nx, st);
}

// sorted: the Morton-ordered records of fps_sort_kernel; N <= 16384
int fps_tmem_launch(const float4 *sorted, int B, int N, int npoint, int tie_ref, int nw, int mp, int32_t *idx, float *nx,
                    cudaStream_t st) {
    if (N <= 4096) return launch_tmem_any<2>(nw, mp, sorted, B, N, npoint, tie_ref, idx, nx, st);
    if (N <= 8192) return launch_tmem_any<4>(nw, mp, sorted, B, N, npoint, tie_ref, idx, nx, st);
    return launch_tmem_any<8>(nw, mp, sorted, B, N, npoint, tie_ref, idx, nx, st);
}

}  // namespace dclr
