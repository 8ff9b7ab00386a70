// Bucketed farthest point sampling, ONE CLOUD SPLIT OVER A CLUSTER OF TWO CTAs.  Same algorithm, arithmetic and tie order
// as fps_tmem.cu (exact bucket pruning, running minima in tensor memory, several exact samples per round); what changes:
//
//   * bucket g of the Z-ordered cloud belongs to CTA g % 2 of the pair (neighbouring buckets alternate, so the update work
//     of a pick is shared), which keeps its points in 96 KB of shared memory and, per bucket, {t lo, t hi, tie ranks} in
//     four TMEM columns (128 columns per CTA).  Two such CTAs fit on one SM (of different clouds at full batch): 16 warps
//     per SM fill the issue slots that one latency-bound CTA leaves empty; a small batch gets two SMs per cloud.
//   * per round every warp publishes its entry (value, key, bound, coordinates) into BOTH CTAs' slot arrays
//     (st.shared::cluster) and arrives on both CTAs' mbarriers (release.cluster); after the acquire-wait all 16 warps of the
//     pair rank the same 16 entries and take the same picks.  16 candidates instead of 8 means more samples per round
//     (tools/fps_multipick_sim.py: 380 -> 333 rounds at MP = 4), for one DSMEM round trip (~215 cycles, B300_MICROARCH) per
//     round in place of a __syncthreads.
#include "common.cuh"
#include "kernels.cuh"
#include "ptx.cuh"
#include "fps_bucket.cuh"
#include <float.h>
#include <limits.h>

namespace dclr {

namespace {

__device__ __forceinline__ float fc_max3(float a, float b, float c) {
    float r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ uint32_t mapa(uint32_t local_addr, uint32_t cta) {       // my shared address -> the peer's
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta));
    return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, int4 v) {
    asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t addr) {                 // local or remote (mapa'd) barrier
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t addr, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra WAIT_%=;\n\t}" ::"r"(addr), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t &a, uint32_t &b, uint32_t &c, uint32_t &d) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_st4(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d)
                 : "memory");
}

}  // namespace

// 8 warps per CTA, PP buckets per warp (local bucket h = i * 8 + warp, global bucket g = 2 h + cluster rank), MP samples per
// round at most; sorted = the Morton-ordered records of fps_sort_kernel.
template <int PP, int MP>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(256, 2)
fps_cluster_kernel(const float4 *__restrict__ sorted, int N, int npoint, int tie_ref, int32_t *__restrict__ idx_out,
                   float *__restrict__ new_xyz_out) {
    extern __shared__ __align__(16) float smem[];
    constexpr int NW = 8, NWT = 16;                                              // warps per CTA / per cloud
    constexpr int CAP = 64 * NW * PP;                                            // records of THIS CTA
    constexpr int WSTRIDE = 64 * NW;
    constexpr int NPAIR = MP * (MP - 1) / 2;
    constexpr int TCOLS = 2 * 4 * PP < 32 ? 32 : 2 * 4 * PP;                     // 128 for 8192 records: {t lo, t hi, ranks, -}
    static_assert(PP <= 32 && MP >= 1 && MP <= NWT && NPAIR <= 32, "shape");
    static_assert((TCOLS & (TCOLS - 1)) == 0 && TCOLS <= 256, "TMEM columns: power of two, two CTAs per SM");
    float *s_x = smem, *s_y = smem + CAP, *s_z = smem + 2 * CAP;
    // warp entries [2][NWT] x {value^sign, ~key, bound, -}, {x, y, z, -}; per warp the first MP winners in order [NW][MP] x same
    int4 *s_slot = reinterpret_cast<int4 *>(smem + 3 * CAP);
    int4 *s_cand = s_slot + 4 * NWT;
    float4 *s_init = reinterpret_cast<float4 *>(s_cand + 2 * NW * MP);           // coordinates of point 0 (either CTA holds it)
    uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_init + 1);
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(s_bar + 1);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.x >> 1;
    uint32_t crank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
    const uint32_t peer = crank ^ 1u;
    sorted += (size_t)b * N;
    idx_out += (size_t)b * npoint;
    if (new_xyz_out) new_xyz_out += (size_t)b * npoint * 3;

    if (warp == 0) ptx::tmem_alloc(s_tmem, TCOLS);
    if (tid == 0) {
        ptx::mbar_init(s_bar, NWT);               // one arrival per warp of the pair and round
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    cluster_sync_all();            // the peer CTA runs and its barrier is initialised: its shared memory may be addressed
    // this thread's state of bucket i: TMEM lane (32 (warp % 4) + lane), columns tcol + 4 i .. + 2
    const uint32_t tmem_base = *s_tmem;
    const uint32_t tcol = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * 4 * PP);

    // ---- stage this CTA's buckets (each lane loads the two records it owns) + per-bucket metadata (lane i <-> bucket i)
    float blx = 0.f, bly = 0.f, blz = 0.f, bhx = 0.f, bhy = 0.f, bhz = 0.f, bm = -1.0f, bm2 = -1.0f;
    const uint32_t wid = crank * NW + (uint32_t)warp;                          // warp id within the cloud
    uint32_t bkey = 0xffffff00u | wid;
    const uint32_t rec0 = (uint32_t)(64 * warp + 2 * lane);                    // local record of pair 0; pair i: + WSTRIDE * i
    const uint32_t ax = ptx::smem_u32(s_x) + rec0 * 4, ay = ptx::smem_u32(s_y) + rec0 * 4,
                   az = ptx::smem_u32(s_z) + rec0 * 4;
    const uint32_t init_a = ptx::smem_u32(s_init);
    constexpr int UNR = PP % 4 == 0 ? 4 : 1;
    for (int i0 = 0; i0 < PP; i0 += UNR) {
        float4 ra[UNR], rb[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const int h = (i0 + u) * NW + warp;                                    // local bucket
            const int gp = 64 * (2 * h + (int)crank) + 2 * lane;                   // global record of my pair
            ra[u] = gp < N ? __ldg(sorted + gp) : make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
            rb[u] = gp + 1 < N ? __ldg(sorted + gp + 1) : make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const int i = i0 + u;
            const int p = (int)rec0 + WSTRIDE * i;
            const int o0 = __float_as_int(ra[u].w), o1 = __float_as_int(rb[u].w);
            const bool v0 = o0 >= 0, v1 = o1 >= 0;
            s_x[p] = ra[u].x; s_y[p] = ra[u].y; s_z[p] = ra[u].z;
            s_x[p + 1] = rb[u].x; s_y[p + 1] = rb[u].y; s_z[p + 1] = rb[u].z;
            const uint32_t k0 = v0 ? (uint32_t)(tie_ref ? fb_rank16(o0) : (uint32_t)o0) & 0xffffu : 0xffffu;
            const uint32_t k1 = v1 ? (uint32_t)(tie_ref ? fb_rank16(o1) : (uint32_t)o1) & 0xffffu : 0xffffu;
            // -1 can never win (reference: best starts at -1, strict >)
            tmem_st4(tcol + 4 * i, __float_as_uint(v0 ? 1e10f : -1.0f), __float_as_uint(v1 ? 1e10f : -1.0f), k0 | (k1 << 16), 0u);
            if (o0 == 0 || o1 == 0) {                                              // point 0 = the first sample: tell both CTAs
                const float4 s = o0 == 0 ? ra[u] : rb[u];
                const int4 v = make_int4(__float_as_int(s.x), __float_as_int(s.y), __float_as_int(s.z), 0);
                st_cluster_v4(mapa(init_a, crank), v);
                st_cluster_v4(mapa(init_a, peer), v);
            }
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
    ptx::tmem_st_wait();
    __syncthreads();
    cluster_sync_all();            // point 0 published to both CTAs

    // candidate pair (pl, pr), pl < pr, tested by this lane: pair index p = pr (pr - 1) / 2 + pl
    int pl = 0, pr = 1;
#pragma unroll
    for (int r = 1; r < MP; ++r)
        if (lane >= r * (r - 1) / 2) { pr = r; pl = lane - r * (r - 1) / 2; }
    const bool pair_lane = lane < NPAIR;
    if (!pair_lane) { pl = 0; pr = MP > 1 ? 1 : 0; }       // (idle lanes read a valid pair; their vote is masked)
    const int4 *cand = s_cand + warp * (2 * MP);

    // picks applied in the coming round: coordinates in registers (unused slots far away: they change nothing)
    float px[MP], py[MP], pz[MP];
#pragma unroll
    for (int q = 0; q < MP; ++q) { px[q] = 1e18f; py[q] = 1e18f; pz[q] = 1e18f; }
    {
        const float4 s = *s_init;
        px[0] = s.x; py[0] = s.y; pz[0] = s.z;
        if (tid == 0 && crank == 0) {
            idx_out[0] = 0;
            if (new_xyz_out) { new_xyz_out[0] = px[0]; new_xyz_out[1] = py[0]; new_xyz_out[2] = pz[0]; }
        }
    }
    const uint32_t bar_l = ptx::smem_u32(s_bar), bar_r = mapa(bar_l, peer);
    int done = 1;                                 // samples selected (and written) so far
    uint32_t ehi = 0u, elo = 0u;                  // this warp's published entry (recomputed only when a bucket changed)
    int ew2 = INT_MIN;
    float ex = 0.f, ey = 0.f, ez = 0.f;
    bool efresh = false;
    for (int round = 0; done < npoint; ++round) {
        // ---- which of this warp's buckets can change?  (lane i tests bucket i against every pick; branch-free)
        unsigned need = 0u;
#pragma unroll
        for (int q = 0; q < MP; ++q) {
            const float dx = fc_max3(blx - px[q], px[q] - bhx, 0.f), dy = fc_max3(bly - py[q], py[q] - bhy, 0.f),
                        dz = fc_max3(blz - pz[q], pz[q] - bhz, 0.f);
            need |= (unsigned)((dx * dx + dy * dy + dz * dz) * 0.9999f < bm);
        }
        unsigned mask = __ballot_sync(0xffffffffu, need != 0u);     // (lanes >= PP keep bm = -1: never set)
        const bool touched = mask != 0u;
        if (touched) {
            uint64_t xo[MP], yo[MP], zo[MP];
#pragma unroll
            for (int q = 0; q < MP; ++q) { xo[q] = fb_pack2(px[q], px[q]); yo[q] = fb_pack2(py[q], py[q]); zo[q] = fb_pack2(pz[q], pz[q]); }
            ptx::tmem_st_wait();                  // the stores of the previous round
            while (mask) {
                // two buckets per iteration (the second repeats the first when only one is left: same values again)
                int iu[2];
                iu[0] = 31 - __clz(mask);
                mask ^= 1u << iu[0];
                iu[1] = mask ? 31 - __clz(mask) : iu[0];
                mask &= ~(1u << iu[1]);
                uint64_t X[2], Y[2], Z[2];
                uint32_t id2[2], tlb[2], thb[2], pad[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    X[u] = fb_lds64(ax + iu[u] * (WSTRIDE * 4)); Y[u] = fb_lds64(ay + iu[u] * (WSTRIDE * 4));
                    Z[u] = fb_lds64(az + iu[u] * (WSTRIDE * 4));
                    tmem_ld4(tcol + 4 * iu[u], tlb[u], thb[u], id2[u], pad[u]);
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
                    b2[u] = __reduce_max_sync(0xffffffffu, holder ? mn[u] : mx[u]);
                    hb[u] = __ballot_sync(0xffffffffu, holder);
                }
#pragma unroll
                for (int u = 0; u < 2; ++u)
                    if (lane == iu[u]) {
                        bm = __int_as_float(bmx[u]);
                        bkey = bmx[u] >= 0 ? kmin[u] : (0xffffff00u | wid);
                        bm2 = __int_as_float(__popc(hb[u]) > 1 ? bmx[u] : b2[u]);
                    }
            }
        }
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
        int4 *slot = s_slot + (round & 1) * (2 * NWT);
        if (lane == 0) {
            // into my CTA's slots and the peer's, then one arrival on each CTA's barrier (release.cluster orders the stores)
            const int4 eA = make_int4((int)ehi, (int)elo, ew2, 0);
            const int4 eB = make_int4(__float_as_int(ex), __float_as_int(ey), __float_as_int(ez), 0);
            slot[2 * wid] = eA;
            slot[2 * wid + 1] = eB;
            const uint32_t ra = mapa(ptx::smem_u32(slot + 2 * wid), peer);
            st_cluster_v4(ra, eA);
            st_cluster_v4(ra + 16, eB);
            mbar_arrive_cluster(bar_l);      // local first: once the peer sees my remote arrival, my local one is counted
            mbar_arrive_cluster(bar_r);
        }
        mbar_wait_cluster(bar_l, (uint32_t)round & 1u);
        // ---- every warp of the pair ranks the 16 entries (value desc, key asc; keys are distinct), first MP in order
        {
            const int me_i = lane < NWT ? lane : 0;
            const int4 eA = slot[2 * me_i], eB = slot[2 * me_i + 1];
            const uint64_t me = ((uint64_t)(uint32_t)eA.x << 32) | (uint32_t)eA.y;
            int c[NWT];
#pragma unroll
            for (int a = 0; a < NWT; ++a) {
                const uint2 o = *reinterpret_cast<const uint2 *>(slot + 2 * a);
                c[a] = (int)((((uint64_t)o.x << 32) | o.y) > me);
            }
#pragma unroll
            for (int s = 1; s < NWT; s <<= 1)
#pragma unroll
                for (int a = 0; a + s < NWT; a += 2 * s) c[a] += c[a + s];
            if (lane < NWT && c[0] < MP) {
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
            uint32_t padw;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(cx[q]), "=r"(cy[q]), "=r"(cz[q]), "=r"(padw)
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
#pragma unroll
        for (int q = 0; q < MP; ++q) {
            const bool on = q < take;
            px[q] = on ? __uint_as_float(cx[q]) : 1e18f; py[q] = on ? __uint_as_float(cy[q]) : 1e18f;
            pz[q] = on ? __uint_as_float(cz[q]) : 1e18f;
        }
        // the record of the picks: lane q of warp 0 of CTA 0 writes sample done + q
        if (crank == 0 && tid < take) {
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
    cluster_sync_all();            // no CTA of the pair leaves while its peer may still address its shared memory
    if (warp == 0) ptx::tmem_dealloc(tmem_base, TCOLS);
}

template <int PP, int MP>
static int launch_cluster(const float4 *sorted, int B, int N, int npoint, int tie_ref, int32_t *idx, float *nx,
                          cudaStream_t st) {
    constexpr size_t smem = (size_t)64 * 8 * PP * 12 + (size_t)(4 * 16 + 2 * 8 * MP) * sizeof(int4) + 16 + 16;
    static_assert(2 * (smem + 1024) <= 233472, "two CTAs per SM");
    auto kern = fps_cluster_kernel<PP, MP>;
    DCLR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<2 * B, 256, smem, st>>>(sorted, N, npoint, tie_ref, idx, nx);
    DCLR_LAUNCH_CHECK("fps_cluster_kernel");
    return DCLR_OK;
}

// sorted: the Morton-ordered records of fps_sort_kernel; 8192 < N <= 16384
int fps_cluster_launch(const float4 *sorted, int B, int N, int npoint, int tie_ref, int mp, int32_t *idx, float *nx,
                       cudaStream_t st) {
    if (mp <= 2) return launch_cluster<16, 2>(sorted, B, N, npoint, tie_ref, idx, nx, st);
    if (mp == 3) return launch_cluster<16, 3>(sorted, B, N, npoint, tie_ref, idx, nx, st);
    if (mp == 4) return launch_cluster<16, 4>(sorted, B, N, npoint, tie_ref, idx, nx, st);
    return launch_cluster<16, 6>(sorted, B, N, npoint, tie_ref, idx, nx, st);
}

}  // namespace dclr
