// Ball query through a uniform grid (large clouds).  Same results as ball_query_kernel (pointops.cu) and therefore as
// the reference's ball_query_kernel_fast (extern/Pointnet2.PyTorch/pointnet2/src/ball_query_gpu.cu:9-45): "the first
// nsample points with d^2 < r^2 in ascending index order", d^2 evaluated with the reference's rounding sequence.  Only
// the set of points that is TESTED changes: instead of all N points per centre, the points of the grid cells the ball
// can touch.
//
//   bq_grid_build_kernel : one CTA per cloud.  Bounding box -> 128 x 128 cells over the two widest axes -> histogram
//                          in shared memory -> exclusive scan -> scatter of (x, y, z, index) records into cell order.
//   bq_grid_query_kernel : one warp per centre.  Conservative cell range of the ball (monotone cell function, radius
//                          inflated by a rounding margin) -> at most a few contiguous record ranges -> exact distance
//                          test -> hits set a bit in a per-warp shared-memory bitmap over point indices; reading the
//                          bitmap in order yields the hits in ascending index order without a sort, for any hit count.
#include "common.cuh"
#include "kernels.cuh"
#include <stdlib.h>
#include <float.h>

namespace dclr {

constexpr int BG_G = 128;                  // cells per axis
constexpr int BG_CELLS = BG_G * BG_G;
constexpr int BG_BUILD_T = 1024;
constexpr int BG_LIST = 64;                // hits ranked from a list instead of the bitmap scan (2 per lane)
constexpr int BG_PARAMS = 8;               // floats per cloud: axis0, axis1 (as int bits), o0, o1, inv0, inv1

__device__ __forceinline__ int bg_cell(float v, float o, float inv) {
    // monotone non-decreasing in v: fp subtract, multiply by a positive number, floor, clamp
    const float f = floorf((v - o) * inv);
    return (int)fminf(fmaxf(f, 0.f), (float)(BG_G - 1));
}

__global__ void __launch_bounds__(BG_BUILD_T, 1)
bq_grid_build_kernel(const float *__restrict__ pts, int stride, int N, float rq, float4 *__restrict__ sorted,
                     int32_t *__restrict__ cellstart, float *__restrict__ params) {
    extern __shared__ int s_cnt[];                       // [BG_CELLS]
    __shared__ float s_red[6][32];
    __shared__ float s_par[BG_PARAMS];
    __shared__ int s_wsum[32];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pts += (size_t)b * N * stride;
    sorted += (size_t)b * N;
    cellstart += (size_t)b * (BG_CELLS + 1);
    params += (size_t)b * BG_PARAMS;

    // ---- bounding box
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int k = tid; k < N; k += BG_BUILD_T) {
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
    for (int e = tid; e < BG_CELLS; e += BG_BUILD_T) s_cnt[e] = 0;
    __syncthreads();
    if (tid == 0) {
        float l[3], h[3], ext[3];
        for (int a = 0; a < 3; ++a) {
            l[a] = FLT_MAX; h[a] = -FLT_MAX;
            for (int w = 0; w < BG_BUILD_T / 32; ++w) { l[a] = fminf(l[a], s_red[a][w]); h[a] = fmaxf(h[a], s_red[3 + a][w]); }
            ext[a] = h[a] - l[a];
        }
        // grid over the two widest axes
        int drop = 0;
        if (ext[1] < ext[drop]) drop = 1;
        if (ext[2] < ext[drop]) drop = 2;
        const int a0 = drop == 0 ? 1 : 0, a1 = drop == 2 ? 1 : 2;
        const float h0 = fmaxf(fmaxf(rq, ext[a0] / BG_G), 1e-30f), h1 = fmaxf(fmaxf(rq, ext[a1] / BG_G), 1e-30f);
        s_par[0] = __int_as_float(a0); s_par[1] = __int_as_float(a1);
        s_par[2] = l[a0]; s_par[3] = l[a1];
        s_par[4] = 1.0f / h0; s_par[5] = 1.0f / h1;
        s_par[6] = 0.f; s_par[7] = 0.f;
    }
    __syncthreads();
    if (tid < BG_PARAMS) params[tid] = s_par[tid];
    const int a0 = __float_as_int(s_par[0]), a1 = __float_as_int(s_par[1]);
    const float o0 = s_par[2], o1 = s_par[3], inv0 = s_par[4], inv1 = s_par[5];

    // ---- histogram
    for (int k = tid; k < N; k += BG_BUILD_T) {
        const float *p = pts + (size_t)k * stride;
        const int c = bg_cell(__ldg(p + a1), o1, inv1) * BG_G + bg_cell(__ldg(p + a0), o0, inv0);
        atomicAdd(&s_cnt[c], 1);
    }
    __syncthreads();
    // ---- exclusive scan of the 16384 counters: 16 consecutive per thread
    constexpr int PER = BG_CELLS / BG_BUILD_T;
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
    for (int i = 0; i < PER; ++i) {
        s_cnt[tid * PER + i] = run;                      // becomes the scatter cursor
        cellstart[tid * PER + i] = run;
        run += v[i];
    }
    if (tid == BG_BUILD_T - 1) cellstart[BG_CELLS] = run;
    __syncthreads();
    // ---- scatter (order inside a cell is irrelevant: the query restores index order through its bitmap)
    for (int k = tid; k < N; k += BG_BUILD_T) {
        const float *p = pts + (size_t)k * stride;
        const float x = __ldg(p), y = __ldg(p + 1), z = __ldg(p + 2);
        const float v0 = a0 == 0 ? x : (a0 == 1 ? y : z), v1 = a1 == 1 ? y : (a1 == 2 ? z : x);
        const int cell = bg_cell(v1, o1, inv1) * BG_G + bg_cell(v0, o0, inv0);
        const int pos = atomicAdd(&s_cnt[cell], 1);
        sorted[pos] = make_float4(x, y, z, __int_as_float(k));
    }
}

struct BgQueryArgs {
    const float *new_xyz;      // [B,M,3]
    const float4 *sorted;      // [B,N] records in cell order
    const int32_t *cellstart;  // [B, BG_CELLS+1]
    const float *params;       // [B, BG_PARAMS]
    int N, M, nr, pad, warps;
    float rq;                  // inflated largest radius used for the cell range
    float r2[2];
    int ns[2];
    int32_t *idx[2];
    int32_t *cnt[2];
};

template <int NR>
__global__ void bq_grid_query_kernel(BgQueryArgs a) {
    extern __shared__ uint32_t s_bm[];          // [warps][NR][words]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.y;
    const int words = (a.N + 31) >> 5;
    uint32_t *bm = s_bm + (size_t)warp * (NR * words + BG_LIST);
    uint32_t *list = bm + NR * words;
    for (int w = lane; w < NR * words; w += 32) bm[w] = 0u;
    __syncwarp();
    const float *par = a.params + (size_t)b * BG_PARAMS;
    const int a0 = __float_as_int(__ldg(par)), a1 = __float_as_int(__ldg(par + 1));
    const float o0 = __ldg(par + 2), o1 = __ldg(par + 3), inv0 = __ldg(par + 4), inv1 = __ldg(par + 5);
    const float4 *rec = a.sorted + (size_t)b * a.N;
    const int32_t *cs = a.cellstart + (size_t)b * (BG_CELLS + 1);

    for (int c = blockIdx.x * a.warps + warp; c < a.M; c += gridDim.x * a.warps) {
        const float *q = a.new_xyz + ((size_t)b * a.M + c) * 3;
        const float cx = __ldg(q), cy = __ldg(q + 1), cz = __ldg(q + 2);
        const float v0 = a0 == 0 ? cx : (a0 == 1 ? cy : cz), v1 = a1 == 1 ? cy : (a1 == 2 ? cz : cx);
        // |c - p| < r (in the kernel's arithmetic) implies p within [v - R, v + R] with a margin for the roundings
        const float R0 = a.rq + fabsf(v0) * 1e-6f, R1 = a.rq + fabsf(v1) * 1e-6f;
        const int x0 = bg_cell(v0 - R0, o0, inv0), x1 = bg_cell(v0 + R0, o0, inv0);
        const int y0 = bg_cell(v1 - R1, o1, inv1), y1 = bg_cell(v1 + R1, o1, inv1);
        int total = 0;                                  // hits of the widest radius so far (warp-uniform)
        for (int y = y0; y <= y1; ++y) {
            const int beg = __ldg(cs + y * BG_G + x0), end = __ldg(cs + y * BG_G + x1 + 1);
            for (int t0 = beg; t0 < end; t0 += 32) {
                const int t = t0 + lane;
                bool any = false;
                uint32_t e = 0u;
                if (t < end) {
                    const float4 p = __ldg(rec + t);
                    // operand order of the reference: (new - x), ball_query_gpu.cu:33
                    const float d2 = sqdist_pn2(cx - p.x, cy - p.y, cz - p.z);
                    const int k = __float_as_int(p.w);
                    e = (uint32_t)k;
#pragma unroll
                    for (int r = 0; r < NR; ++r)
                        if (d2 < a.r2[r]) {
                            atomicOr(&bm[r * words + (k >> 5)], 1u << (k & 31));
                            e |= 0x80000000u >> r;      // membership flags ride in the top bits (N < 2^30)
                            any = true;
                        }
                }
                // short hit list next to the bitmap: sparse neighbourhoods are ranked from it (below)
                const unsigned hm = __ballot_sync(0xffffffffu, any);
                if (hm) {
                    const int pos = total + __popc(hm & ((1u << lane) - 1u));
                    if (any && pos < BG_LIST) list[pos] = e;
                    total += __popc(hm);
                }
            }
        }
        __syncwarp();
        if (total <= BG_LIST) {
            // ---- sparse centre (the usual case on LiDAR scans away from the sensor): the place of a hit in the
            //      ascending-index order is the number of listed hits of the same radius with a smaller index
            const uint32_t e0 = lane < total ? list[lane] : 0u, e1 = lane + 32 < total ? list[lane + 32] : 0u;
            const uint32_t i0 = e0 & 0x3fffffffu, i1 = e1 & 0x3fffffffu;
            int rk0[NR], rk1[NR];
#pragma unroll
            for (int r = 0; r < NR; ++r) rk0[r] = rk1[r] = 0;
            for (int j = 0; j < total; ++j) {
                const uint32_t ej = list[j];
                const uint32_t ij = ej & 0x3fffffffu;
#pragma unroll
                for (int r = 0; r < NR; ++r) {
                    const bool fj = (ej << r) >> 31;
                    rk0[r] += (fj && ij < i0) ? 1 : 0;
                    rk1[r] += (fj && ij < i1) ? 1 : 0;
                }
            }
#pragma unroll
            for (int r = 0; r < NR; ++r) {
                int32_t *row = a.idx[r] + ((size_t)b * a.M + c) * a.ns[r];
                const int ns = a.ns[r];
                const bool f0 = (e0 << r) >> 31, f1 = (e1 << r) >> 31;
                if (f0) { bm[r * words + (i0 >> 5)] = 0u; if (rk0[r] < ns) row[rk0[r]] = (int32_t)i0; }
                if (f1) { bm[r * words + (i1 >> 5)] = 0u; if (rk1[r] < ns) row[rk1[r]] = (int32_t)i1; }
                const int cnt = __popc(__ballot_sync(0xffffffffu, f0)) + __popc(__ballot_sync(0xffffffffu, f1));
                const int n = min(cnt, ns);
                if (a.cnt[r] && lane == 0) a.cnt[r][(size_t)b * a.M + c] = n;
                if (a.pad) {
                    const uint32_t m0 = f0 ? i0 : 0xffffffffu, m1 = f1 ? i1 : 0xffffffffu;
                    const uint32_t first = __reduce_min_sync(0xffffffffu, min(m0, m1));
                    const int fill = n > 0 ? (int)first : 0;
                    for (int p2 = n + lane; p2 < ns; p2 += 32) row[p2] = fill;
                }
            }
            __syncwarp();
            continue;
        }
        // read the bitmaps in index order (and clear them for the next centre)
#pragma unroll
        for (int r = 0; r < NR; ++r) {
            int32_t *row = a.idx[r] + ((size_t)b * a.M + c) * a.ns[r];
            const int ns = a.ns[r];
            int cnt = 0, first = 0;
            for (int w0 = 0; w0 < words; w0 += 32) {
                const int w = w0 + lane;
                uint32_t word = 0u;
                if (w < words) { word = bm[r * words + w]; if (word) bm[r * words + w] = 0u; }
                if (__ballot_sync(0xffffffffu, word != 0u) == 0u) continue;
                const int mine = __popc(word);
                int incl = mine;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int o = __shfl_up_sync(0xffffffffu, incl, d);
                    if (lane >= d) incl += o;
                }
                int pos = cnt + incl - mine;
                if (cnt == 0) {
                    // first hit overall = lowest set bit of the lowest lane with a non-empty word
                    const unsigned nz = __ballot_sync(0xffffffffu, word != 0u);
                    const int src = __ffs(nz) - 1;
                    const uint32_t fw = __shfl_sync(0xffffffffu, word, src);
                    first = (w0 + src) * 32 + __ffs(fw) - 1;
                }
                while (word && pos < ns) {
                    const int bit = __ffs(word) - 1;
                    row[pos++] = w * 32 + bit;
                    word &= word - 1;
                }
                cnt += __shfl_sync(0xffffffffu, incl, 31);
            }
            const int n = min(cnt, ns);
            if (a.cnt[r] && lane == 0) a.cnt[r][(size_t)b * a.M + c] = n;
            if (a.pad) {
                const int fill = n > 0 ? first : 0;
                for (int p2 = n + lane; p2 < ns; p2 += 32) row[p2] = fill;
            }
        }
        __syncwarp();
    }
}

// Returns DCLR_OK and sets *used = 1 when the grid path ran; *used = 0 when the shape is outside what it is built for
// (caller falls back to the scan kernel).  Workspace: sorted [B*N] float4, cellstart [B*(BG_CELLS+1)] int32,
// params [B*BG_PARAMS] float.
size_t ball_grid_workspace_bytes(int B, int N) {
    return (size_t)B * N * sizeof(float4) + (size_t)B * (BG_CELLS + 1) * sizeof(int32_t) +
           (size_t)B * BG_PARAMS * sizeof(float) + 512;
}

int ball_grid_launch(const float *new_xyz, const float *pts, int stride, int B, int N, int M, int nr,
                     const float *radii, const int *ns, int32_t *const *idx, int32_t *const *cnt, int pad, void *ws,
                     cudaStream_t st, int *used, int phase) {
    *used = 0;
    const int words = (N + 31) / 32;
    int warps = (int)((200u * 1024u) / (((size_t)nr * words + BG_LIST) * 4));
    if (warps > 8) warps = 8;
    const char *env_min = getenv("DCLR_BQ_GRID_MIN");            // tuning knob (tools): smallest cloud that takes the grid
    const int min_n = env_min ? atoi(env_min) : BG_MIN_N;   // 1024: ModelNet-shaped batch 1.83 -> 1.66 ms (2048 before)
    if (N < min_n || warps < 1 || B > 65535) return DCLR_OK;
    float rmax = 0.f;
    for (int r = 0; r < nr; ++r) rmax = radii[r] > rmax ? radii[r] : rmax;
    if (!(rmax > 0.f)) return DCLR_OK;
    const float rq = rmax * 1.001f;
    char *w = (char *)ws;
    float4 *sorted = (float4 *)w;
    int32_t *cellstart = (int32_t *)(w + (size_t)B * N * sizeof(float4));
    float *params = (float *)(w + (size_t)B * N * sizeof(float4) + (((size_t)B * (BG_CELLS + 1) * 4 + 255) / 256) * 256);
    // the opt-in is per DEVICE (a process may run models on several GPUs): one flag per device ordinal, idempotent set,
    // benign if raced
    static bool attr_done[64] = {false};
    int dev = 0;
    DCLR_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_done[dev]) {
        DCLR_CUDA(cudaFuncSetAttribute(bq_grid_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       BG_CELLS * (int)sizeof(int)));
        DCLR_CUDA(cudaFuncSetAttribute(bq_grid_query_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        DCLR_CUDA(cudaFuncSetAttribute(bq_grid_query_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        if (dev >= 0 && dev < 64) attr_done[dev] = true;
    }
    if (phase != BG_PHASE_QUERY) {
        bq_grid_build_kernel<<<B, BG_BUILD_T, BG_CELLS * sizeof(int), st>>>(pts, stride, N, rq, sorted, cellstart, params);
        DCLR_LAUNCH_CHECK("bq_grid_build_kernel");
    }
    *used = 1;
    if (phase == BG_PHASE_BUILD) return DCLR_OK;
    BgQueryArgs a;
    a.new_xyz = new_xyz; a.sorted = sorted; a.cellstart = cellstart; a.params = params;
    a.N = N; a.M = M; a.nr = nr; a.pad = pad; a.warps = warps; a.rq = rq;
    for (int r = 0; r < 2; ++r) {
        const int rr = r < nr ? r : 0;
        a.r2[r] = radii[rr] * radii[rr];            // float radius2 = radius * radius  (ball_query_gpu.cu:23)
        a.ns[r] = ns[rr]; a.idx[r] = idx[rr]; a.cnt[r] = cnt ? cnt[rr] : nullptr;
    }
    const size_t smem = (size_t)warps * ((size_t)nr * words + BG_LIST) * 4;
    // enough CTAs per cloud to fill the GPU, each warp looping over centres
    int per_cloud = div_up(M, warps);
    const int want = div_up(148 * 4, B);
    if (per_cloud > want) per_cloud = want < 1 ? 1 : want;
    dim3 grid(per_cloud, B);
    if (nr == 2) bq_grid_query_kernel<2><<<grid, warps * 32, smem, st>>>(a);
    else bq_grid_query_kernel<1><<<grid, warps * 32, smem, st>>>(a);
    DCLR_LAUNCH_CHECK("bq_grid_query_kernel");
    *used = 1;
    return DCLR_OK;
}

}  // namespace dclr
