// gather / group / ball query / kNN for sm_100a.
//
// ball query: replaces ball_query_kernel_fast (extern/Pointnet2.PyTorch/pointnet2/src/ball_query_gpu.cu:9-45).  The
// reference runs one thread per centre over all N points with strided scalar loads; here a warp owns CPW centres,
// its lanes take 32 consecutive points of a tile that the TMA engine (cp.async.bulk, mbarrier-tracked, double
// buffered) staged into shared memory, and hits are compacted in index order with ballot/popc, so the result is the
// reference's "first nsample hits in ascending index order".  Up to two radii share one distance pass.
//
// kNN: replaces torch_cluster.knn as called at deepclr/models/deepclr.py:164-166; a warp owns one query, candidates
// are packed as (float bits of d^2 << 32 | index) so a u64 minimum is "nearest, lowest index on ties".
#include "common.cuh"
#include "kernels.cuh"
#include "ptx.cuh"

namespace dclr {

// ------------------------------------------------------------------ gather / group (compat ops)
__global__ void gather_kernel(const float *__restrict__ feat, const int32_t *__restrict__ idx, int C, int N, int M,
                              float *__restrict__ out) {
    int b = blockIdx.z, c = blockIdx.y, j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= M) return;
    out[((size_t)b * C + c) * M + j] = __ldg(feat + ((size_t)b * C + c) * N + __ldg(idx + (size_t)b * M + j));
}

__global__ void group_kernel(const float *__restrict__ feat, const int32_t *__restrict__ idx, int C, int N, int M,
                             int ns, float *__restrict__ out) {
    int b = blockIdx.z, c = blockIdx.y;
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;   // over M*ns
    if (e >= (size_t)M * ns) return;
    out[((size_t)b * C + c) * M * ns + e] = __ldg(feat + ((size_t)b * C + c) * N + __ldg(idx + (size_t)b * M * ns + e));
}

// backward of the two: scatter-add of the upstream gradient into the source positions.  Same arithmetic as the
// reference's gather_points_grad_kernel_fast (sampling_gpu.cu:46-63) and group_points_grad_kernel_fast
// (group_points_gpu.cu:8-25): fp32 atomicAdd per element, so equal to them up to the order of the additions.
__global__ void gather_grad_kernel(const float *__restrict__ grad_out, const int32_t *__restrict__ idx, int C, int N,
                                   int M, float *__restrict__ grad_points) {
    int b = blockIdx.z, c = blockIdx.y, j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= M) return;
    atomicAdd(grad_points + ((size_t)b * C + c) * N + __ldg(idx + (size_t)b * M + j),
              __ldg(grad_out + ((size_t)b * C + c) * M + j));
}

__global__ void group_grad_kernel(const float *__restrict__ grad_out, const int32_t *__restrict__ idx, int C, int N,
                                  int M, int ns, float *__restrict__ grad_points) {
    int b = blockIdx.z, c = blockIdx.y;
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;   // over M*ns
    if (e >= (size_t)M * ns) return;
    atomicAdd(grad_points + ((size_t)b * C + c) * N + __ldg(idx + (size_t)b * M * ns + e),
              __ldg(grad_out + ((size_t)b * C + c) * M * ns + e));
}

// ------------------------------------------------------------------ ball query
constexpr int BQ_TILE = 2048;     // points per stage
constexpr int BQ_WARPS = 8;

struct BqArgs {
    const float *new_xyz;   // [B,M,3]
    const float *pts;       // [B,N,stride]
    int stride, N, M;
    float r2[2];
    int ns[2];
    int32_t *idx[2];        // [B,M,ns]
    int32_t *cnt[2];        // [B,M] nullable
    int pad;
};

template <int NR, int CPW>
__global__ void __launch_bounds__(BQ_WARPS * 32)
ball_query_kernel(BqArgs a) {
    extern __shared__ __align__(128) float bq_smem[];
    __shared__ __align__(8) uint64_t bars[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.y;
    const int N = a.N, M = a.M, stride = a.stride;
    const float *pts = a.pts + (size_t)b * N * stride;
    const bool bulk = (stride == 4 || stride == 3) && ((reinterpret_cast<uintptr_t>(pts) & 15) == 0) &&
                      (((size_t)N * stride * 4) % 16 == 0);
    const int ss = bulk ? stride : 3;                 // point stride in shared memory
    float *buf0 = bq_smem, *buf1 = bq_smem + BQ_TILE * 4;
    const int ntiles = (N + BQ_TILE - 1) / BQ_TILE;

    if (tid == 0) {
        ptx::mbar_init(&bars[0], 1);
        ptx::mbar_init(&bars[1], 1);
        ptx::fence_mbar_init();
    }
    __syncthreads();
    auto issue = [&](int t) {   // one thread: stage tile t
        int cntp = min(BQ_TILE, N - t * BQ_TILE);
        uint32_t bytes = (uint32_t)cntp * stride * 4;
        ptx::mbar_expect_tx(&bars[t & 1], bytes);
        ptx::bulk_g2s((t & 1) ? buf1 : buf0, pts + (size_t)t * BQ_TILE * stride, bytes, &bars[t & 1]);
    };
    if (bulk && tid == 0) issue(0);

    // centres of this warp
    const int c0 = (blockIdx.x * BQ_WARPS + warp) * CPW;
    float cx[CPW], cy[CPW], cz[CPW];
    int cnt[CPW][NR], first[CPW][NR];
#pragma unroll
    for (int ci = 0; ci < CPW; ++ci) {
        int c = c0 + ci;
        const float *q = a.new_xyz + ((size_t)b * M + (c < M ? c : 0)) * 3;
        cx[ci] = __ldg(q); cy[ci] = __ldg(q + 1); cz[ci] = __ldg(q + 2);
#pragma unroll
        for (int r = 0; r < NR; ++r) { cnt[ci][r] = (c < M) ? 0 : 0x40000000; first[ci][r] = 0; }
    }
    float r2max = a.r2[0];
    if (NR > 1) r2max = fmaxf(r2max, a.r2[1]);
    const unsigned lt_mask = (1u << lane) - 1u;

    for (int t = 0; t < ntiles; ++t) {
        const int base = t * BQ_TILE;
        const int tile_cnt = min(BQ_TILE, N - base);
        float *buf = (t & 1) ? buf1 : buf0;
        if (bulk) {
            if (tid == 0 && t + 1 < ntiles) issue(t + 1);
            ptx::mbar_wait(&bars[t & 1], (t >> 1) & 1);
        } else {
            for (int k = tid; k < tile_cnt; k += BQ_WARPS * 32) {
                const float *p = pts + (size_t)(base + k) * stride;
                buf[k * 3 + 0] = __ldg(p); buf[k * 3 + 1] = __ldg(p + 1); buf[k * 3 + 2] = __ldg(p + 2);
            }
            __syncthreads();
        }
        bool done = true;
#pragma unroll
        for (int ci = 0; ci < CPW; ++ci)
#pragma unroll
            for (int r = 0; r < NR; ++r) done = done && (cnt[ci][r] >= a.ns[r]);
        if (!done) {
            for (int s = 0; s < tile_cnt; s += 32) {
                const int kk = s + lane;
                const bool valid = kk < tile_cnt;
                float px = 0.f, py = 0.f, pz = 0.f;
                if (valid) {
                    if (ss == 4) {
                        float4 v = *reinterpret_cast<const float4 *>(buf + kk * 4);
                        px = v.x; py = v.y; pz = v.z;
                    } else {
                        px = buf[kk * 3]; py = buf[kk * 3 + 1]; pz = buf[kk * 3 + 2];
                    }
                }
#pragma unroll
                for (int ci = 0; ci < CPW; ++ci) {
                    // operand order of the reference: (new - x), ball_query_gpu.cu:33
                    const float d2 = sqdist_pn2(cx[ci] - px, cy[ci] - py, cz[ci] - pz);
                    const bool h = valid && (d2 < r2max);
                    const unsigned any = __ballot_sync(0xffffffffu, h);
                    if (any) {
#pragma unroll
                        for (int r = 0; r < NR; ++r) {
                            const bool hr = h && (d2 < a.r2[r]);
                            const unsigned mr = (NR == 1) ? any : __ballot_sync(0xffffffffu, hr);
                            if (mr) {
                                const int pos = cnt[ci][r] + __popc(mr & lt_mask);
                                if (hr && pos < a.ns[r])
                                    a.idx[r][((size_t)b * M + c0 + ci) * a.ns[r] + pos] = base + kk;
                                if (cnt[ci][r] == 0) first[ci][r] = base + s + __ffs(mr) - 1;
                                cnt[ci][r] += __popc(mr);
                            }
                        }
                    }
                }
            }
        }
        __syncthreads();   // all warps finished with this buffer before it is refilled
    }

    // epilogue: counts + padding (ball_query_gpu.cu:35-39: every slot is pre-filled with the first hit; empty ball =
    // caller's zeros, pointnet2_utils.py:218)
#pragma unroll
    for (int ci = 0; ci < CPW; ++ci) {
        const int c = c0 + ci;
        if (c >= M) continue;
#pragma unroll
        for (int r = 0; r < NR; ++r) {
            const int n = min(cnt[ci][r], a.ns[r]);
            if (a.cnt[r] && lane == 0) a.cnt[r][(size_t)b * M + c] = n;
            if (a.pad) {
                const int fill = n > 0 ? first[ci][r] : 0;
                int32_t *row = a.idx[r] + ((size_t)b * M + c) * a.ns[r];
                for (int p = n + lane; p < a.ns[r]; p += 32) row[p] = fill;
            }
        }
    }
}

template <int NR>
static int bq_launch(const BqArgs &a, int B, cudaStream_t st) {
    // fewer centres per warp when the grid would not fill the GPU
    const size_t smem = (size_t)2 * BQ_TILE * 4 * sizeof(float);
    long ctas8 = (long)div_up(a.M, 8 * BQ_WARPS) * B;
    if (ctas8 >= 148) {
        auto k = ball_query_kernel<NR, 8>;
        DCLR_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<dim3(div_up(a.M, 8 * BQ_WARPS), B), BQ_WARPS * 32, smem, st>>>(a);
    } else {
        auto k = ball_query_kernel<NR, 2>;
        DCLR_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<dim3(div_up(a.M, 2 * BQ_WARPS), B), BQ_WARPS * 32, smem, st>>>(a);
    }
    DCLR_LAUNCH_CHECK("ball_query_kernel");
    return DCLR_OK;
}

int ball_query_launch(const float *new_xyz, const float *pts, int stride, int B, int N, int M, int nr,
                      const float *radii, const int *ns, int32_t *const *idx, int32_t *const *cnt, int pad,
                      cudaStream_t st) {
    BqArgs a;
    a.new_xyz = new_xyz; a.pts = pts; a.stride = stride; a.N = N; a.M = M; a.pad = pad;
    for (int r = 0; r < 2; ++r) {
        int rr = r < nr ? r : 0;
        a.r2[r] = radii[rr] * radii[rr];            // float radius2 = radius * radius  (ball_query_gpu.cu:23)
        a.ns[r] = ns[rr]; a.idx[r] = idx[rr]; a.cnt[r] = cnt ? cnt[rr] : nullptr;
    }
    return nr == 2 ? bq_launch<2>(a, B, st) : bq_launch<1>(a, B, st);
}

// ------------------------------------------------------------------ kNN
// General kernel (any Nx, k <= 128).  A warp owns KNN_QPW queries; the candidates pass through shared memory in tiles of
// 1024 (structure of arrays, so a lane reads two neighbouring candidates per 8-byte load and evaluates them with one
// packed f32x2 instruction per step, each half rounded exactly like the scalar expression).  The k best so far live in
// shared memory as sorted keys (d bits << 32 | index: "ascending d^2, ties -> lower index", the oracle's order).  After
// the first tile the bar is the current k-th distance: a tile contributes only the candidates at or under it -- on
// unordered clouds ~k ln(Nx / k) of them over ALL later tiles -- so most tiles cost their 32 distances per lane, one
// count and a vote.  The few survivors are compacted and merged with the current winners by k rounds of a warp minimum
// over at most four keys per lane.  A tile with too many survivors (the first one; later ones while the bar is loose, as on
// clouds stored in scan order) first tightens the bar with its own k-th smallest lane minimum; only if even that leaves too
// many (masses of ties) the k minima are extracted from all of its keys.  16384 x 16384 x 8 clouds, k = 20: 23 ms with the
// per-tile full extraction only.
constexpr int KNN_WARPS = 8;
constexpr int KNN_QPW = 2;                  // queries per warp (share a staged tile)
constexpr int KNN_KPL = 32;                 // keys per lane per tile
constexpr int KNN_TILE = 32 * KNN_KPL;      // 1024 candidates per tile
constexpr int KNN_MAXK = 128;
constexpr int KNN_CAP = 128;                // merge capacity: current winners + survivors of a tile

__device__ __forceinline__ uint64_t warp_min_u64(uint64_t v) {
    uint32_t hi = (uint32_t)(v >> 32);
    uint32_t mhi = __reduce_min_sync(0xffffffffu, hi);
    uint32_t lo = (hi == mhi) ? (uint32_t)v : 0xffffffffu;
    uint32_t mlo = __reduce_min_sync(0xffffffffu, lo);
    return ((uint64_t)mhi << 32) | mlo;
}

__global__ void __launch_bounds__(KNN_WARPS * 32, 3)
knn_kernel(const float *__restrict__ x, const float *__restrict__ y, int Nx, int Ny, int k,
           int64_t *__restrict__ row_out, int64_t *__restrict__ col_out, int32_t *__restrict__ col_local,
           float4 *__restrict__ nbr4) {
    __shared__ __align__(8) float s_x[KNN_TILE], s_y[KNN_TILE], s_z[KNN_TILE];
    __shared__ uint64_t s_win[KNN_WARPS][KNN_QPW][KNN_MAXK];
    __shared__ uint64_t s_cand[KNN_WARPS][KNN_CAP];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.y;
    const int q0 = (blockIdx.x * KNN_WARPS + warp) * KNN_QPW;
    const float *xb = x + (size_t)b * Nx * 3;
    const uint64_t KMAX = ~0ull;
    float qx[KNN_QPW], qy[KNN_QPW], qz[KNN_QPW];
    uint32_t bar[KNN_QPW];                                          // bits of the k-th best distance so far
#pragma unroll
    for (int qi = 0; qi < KNN_QPW; ++qi) {
        const size_t gy = (size_t)b * Ny + min(q0 + qi, Ny - 1);
        qx[qi] = __ldg(y + gy * 3); qy[qi] = __ldg(y + gy * 3 + 1); qz[qi] = __ldg(y + gy * 3 + 2);
        bar[qi] = 0xffffffffu;
    }
    const int ntiles = (Nx + KNN_TILE - 1) / KNN_TILE;
    for (int t = 0; t < ntiles; ++t) {
        const int base = t * KNN_TILE;
        const int tile_cnt = min(KNN_TILE, Nx - base);
        __syncthreads();
        for (int j = tid; j < KNN_TILE; j += KNN_WARPS * 32) {
            const bool ok = j < tile_cnt;
            const float *c = xb + (size_t)(base + (ok ? j : 0)) * 3;
            s_x[j] = __ldg(c); s_y[j] = __ldg(c + 1); s_z[j] = __ldg(c + 2);
        }
        __syncthreads();
#pragma unroll
        for (int qi = 0; qi < KNN_QPW; ++qi) {
            if (q0 + qi >= Ny) continue;                            // warp-uniform
            // d[2 i], d[2 i + 1] = tile candidates 2 (i * 32 + lane), + 1.  torch-cluster: tmp_dist += (x - y)^2 per
            // dimension, x = candidate, y = query
            float d[KNN_KPL];
            {
                uint64_t qxp, qyp, qzp;
                asm("mov.b64 %0, {%1, %1};" : "=l"(qxp) : "f"(qx[qi]));
                asm("mov.b64 %0, {%1, %1};" : "=l"(qyp) : "f"(qy[qi]));
                asm("mov.b64 %0, {%1, %1};" : "=l"(qzp) : "f"(qz[qi]));
#pragma unroll
                for (int i = 0; i < KNN_KPL / 2; ++i) {
                    const int j = 2 * (i * 32 + lane);
                    uint64_t dx, dy, dz, acc;
                    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(dx) : "l"(*reinterpret_cast<const uint64_t *>(s_x + j)), "l"(qxp));
                    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(dy) : "l"(*reinterpret_cast<const uint64_t *>(s_y + j)), "l"(qyp));
                    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(dz) : "l"(*reinterpret_cast<const uint64_t *>(s_z + j)), "l"(qzp));
                    asm("mul.rn.f32x2 %0, %1, %1;" : "=l"(acc) : "l"(dx));
                    asm("fma.rn.f32x2 %0, %1, %1, %2;" : "=l"(acc) : "l"(dy), "l"(acc));
                    asm("fma.rn.f32x2 %0, %1, %1, %2;" : "=l"(acc) : "l"(dz), "l"(acc));
                    asm("mov.b64 {%0, %1}, %2;" : "=f"(d[2 * i]), "=f"(d[2 * i + 1]) : "l"(acc));
                }
            }
            // survivors: existing candidates at or under the bar (d >= 0: its bit image is ordered)
            uint32_t bq = bar[qi];
            int cnt = 0;
#pragma unroll
            for (int i = 0; i < KNN_KPL; ++i) {
                const int j = 2 * ((i >> 1) * 32 + lane) + (i & 1);
                cnt += (j < tile_cnt && __float_as_uint(d[i]) <= bq) ? 1 : 0;
            }
            int total = __reduce_add_sync(0xffffffffu, cnt);
            if (total == 0) continue;                               // warp-uniform: the usual case after a few tiles
            uint64_t *win = s_win[warp][qi];
            const int nw = t > 0 ? k : 0;                           // winners carried over
            if (total + nw > KNN_CAP) {
                // too many for the merge (always in the first tile; later when the cloud is ordered in space and the
                // bar is still loose): the tile itself bounds the answer.  T = k-th smallest of the 32 per-lane minima:
                // at least k of its candidates are <= T, so nothing above min(bar, T) can be among the k best.
                uint32_t v = 0xffffffffu, T = 0xffffffffu;
#pragma unroll
                for (int i = 0; i < KNN_KPL; ++i) {
                    const int j = 2 * ((i >> 1) * 32 + lane) + (i & 1);
                    if (j < tile_cnt) v = min(v, __float_as_uint(d[i]));
                }
                for (int r = 0; r < k; ++r) {
                    T = __reduce_min_sync(0xffffffffu, v);
                    if (v == T) v = 0xffffffffu;
                }
                bq = min(bq, T);
                cnt = 0;
#pragma unroll
                for (int i = 0; i < KNN_KPL; ++i) {
                    const int j = 2 * ((i >> 1) * 32 + lane) + (i & 1);
                    cnt += (j < tile_cnt && __float_as_uint(d[i]) <= bq) ? 1 : 0;
                }
                total = __reduce_add_sync(0xffffffffu, cnt);
            }
            if (total + nw <= KNN_CAP) {
                // ---- merge: compact the survivors, then the k minima of (winners + survivors), <= 4 keys per lane
                int off = cnt;
#pragma unroll
                for (int sft = 1; sft < 32; sft <<= 1) {
                    const int o = __shfl_up_sync(0xffffffffu, off, sft);
                    if (lane >= sft) off += o;
                }
                off -= cnt;
                uint64_t *cand = s_cand[warp];
#pragma unroll
                for (int i = 0; i < KNN_KPL; ++i) {
                    const int j = 2 * ((i >> 1) * 32 + lane) + (i & 1);
                    if (j < tile_cnt && __float_as_uint(d[i]) <= bq)
                        cand[off++] = ((uint64_t)__float_as_uint(d[i]) << 32) | (uint32_t)(base + j);
                }
                __syncwarp();
                uint64_t c[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int e = lane + 32 * u;
                    c[u] = e < nw ? win[e] : (e - nw < total ? cand[e - nw] : KMAX);
                }
                __syncwarp();
                uint64_t lmin = min(min(c[0], c[1]), min(c[2], c[3]));
                for (int r = 0; r < k; ++r) {
                    const uint64_t w = warp_min_u64(lmin);
                    if (lane == 0) win[r] = w;
                    if (lmin == w) {                                 // exactly one lane (indices are unique)
#pragma unroll
                        for (int u = 0; u < 4; ++u) c[u] = c[u] == w ? KMAX : c[u];
                        lmin = min(min(c[0], c[1]), min(c[2], c[3]));
                    }
                }
            } else {
                // ---- all keys of the tile + the winners carried over from earlier tiles
                uint64_t keys[KNN_KPL + 4];
#pragma unroll
                for (int i = 0; i < KNN_KPL; ++i) {
                    const int j = 2 * ((i >> 1) * 32 + lane) + (i & 1);
                    keys[i] = j < tile_cnt ? ((uint64_t)__float_as_uint(d[i]) << 32) | (uint32_t)(base + j) : KMAX;
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int w = lane + 32 * u;
                    keys[KNN_KPL + u] = (t > 0 && w < k) ? win[w] : KMAX;
                }
                __syncwarp();
                uint64_t lmin = KMAX;
#pragma unroll
                for (int i = 0; i < KNN_KPL + 4; ++i) lmin = keys[i] < lmin ? keys[i] : lmin;
                for (int r = 0; r < k; ++r) {
                    const uint64_t w = warp_min_u64(lmin);
                    if (lane == 0) win[r] = w;
                    if (lmin == w) {                                 // exactly one lane (indices are unique)
                        uint64_t nm = KMAX;
#pragma unroll
                        for (int i = 0; i < KNN_KPL + 4; ++i) {
                            uint64_t v = keys[i];
                            v = (v == w) ? KMAX : v;
                            keys[i] = v;
                            nm = v < nm ? v : nm;
                        }
                        lmin = nm;
                    }
                }
            }
            __syncwarp();
            bar[qi] = (uint32_t)(win[k - 1] >> 32);                 // 0xffffffff while fewer than k candidates were seen
        }
    }
#pragma unroll
    for (int qi = 0; qi < KNN_QPW; ++qi) {
        if (q0 + qi >= Ny) continue;
        const size_t gy = (size_t)b * Ny + q0 + qi;
        for (int r = lane; r < k; r += 32) {
            const uint32_t ci = (uint32_t)s_win[warp][qi][r];
            const size_t o = gy * k + r;
            if (row_out) row_out[o] = (int64_t)gy;
            if (col_out) col_out[o] = (int64_t)((size_t)b * Nx + ci);
            if (col_local) col_local[o] = (int32_t)ci;
            if (nbr4) {
                const float *c = xb + (size_t)ci * 3;
                nbr4[o] = make_float4(__ldg(c) - qx[qi], __ldg(c + 1) - qy[qi], __ldg(c + 2) - qz[qi],
                                      __int_as_float((int)ci));
            }
        }
    }
}

// Fast path for Nx <= 1024, k <= 32 (the model's shapes: queries and candidates are the FPS samples).  A warp owns
// a query; lane l holds the squared distances to candidates l, l+32, ... in registers.  Instead of extracting the k
// minima from all Nx keys, a threshold T = k-th smallest of the 32 per-lane minima is taken first: at least k
// candidates have d <= T (one per contributing lane), typically ~1.5k do.  Those few are compacted to shared memory
// as (d bits << 32 | index) keys and the k ordered minima are extracted from at most 4 keys per lane.  Ordering by
// the packed key is "ascending d^2, ties -> lower index", the same order as knn_kernel / the oracle.
constexpr int KNNF_WARPS = 8;
constexpr int KNNF_QPW = 4;                 // queries per warp (share the staged candidates)
constexpr int KNNF_CAP = 128;               // compacted candidates per query
#ifndef DCLR_KNN_TWOMIN_K
#define DCLR_KNN_TWOMIN_K 24
#endif
constexpr int KNNF_TWOMIN_K = DCLR_KNN_TWOMIN_K;   // above this k the threshold uses two minima per lane

template <int KPL>
__global__ void __launch_bounds__(KNNF_WARPS * 32)
knn_fast_kernel(const float *__restrict__ x, const float *__restrict__ y, int Nx, int Ny, int k,
                int64_t *__restrict__ row_out, int64_t *__restrict__ col_out, int32_t *__restrict__ col_local,
                float4 *__restrict__ nbr4) {
    __shared__ __align__(8) float s_x[KPL * 32], s_y[KPL * 32], s_z[KPL * 32];
    __shared__ uint64_t s_cand[KNNF_WARPS][KNNF_CAP];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.y;
    const float *xb = x + (size_t)b * Nx * 3;
    // candidates that do not exist sit far away (d = 3e36: finite, never among the k <= Nx nearest) -- no bounds test in
    // the distance loop
    for (int j = tid; j < KPL * 32; j += KNNF_WARPS * 32) {
        const bool ok = j < Nx;
        s_x[j] = ok ? __ldg(xb + j * 3) : 1e18f;
        s_y[j] = ok ? __ldg(xb + j * 3 + 1) : 1e18f;
        s_z[j] = ok ? __ldg(xb + j * 3 + 2) : 1e18f;
    }
    __syncthreads();
    static_assert(KPL % 2 == 0, "candidate pairs");
    const uint64_t KMAX = ~0ull;
    const unsigned lt_mask = (1u << lane) - 1u;
    for (int qi = 0; qi < KNNF_QPW; ++qi) {
        const int q = (blockIdx.x * KNNF_WARPS + warp) * KNNF_QPW + qi;
        if (q >= Ny) break;                                     // warp-uniform
        const size_t gy = (size_t)b * Ny + q;
        const float qx = __ldg(y + gy * 3), qy = __ldg(y + gy * 3 + 1), qz = __ldg(y + gy * 3 + 2);
        // d[2 i], d[2 i + 1] = candidates 2 (i * 32 + lane), + 1: two per packed instruction, each half rounded exactly like
        // the scalar expression.  torch-cluster: tmp_dist += (x - y)^2 per dimension, x = candidate, y = query
        float d[KPL];
        float lmin = INFINITY;
        {
            uint64_t qxp, qyp, qzp;
            asm("mov.b64 %0, {%1, %1};" : "=l"(qxp) : "f"(qx));
            asm("mov.b64 %0, {%1, %1};" : "=l"(qyp) : "f"(qy));
            asm("mov.b64 %0, {%1, %1};" : "=l"(qzp) : "f"(qz));
#pragma unroll
            for (int i = 0; i < KPL / 2; ++i) {
                const int j = 2 * (i * 32 + lane);            // one 8-byte shared load per coordinate and candidate pair
                uint64_t dx, dy, dz, acc;
                asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(dx) : "l"(*reinterpret_cast<const uint64_t *>(s_x + j)), "l"(qxp));
                asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(dy) : "l"(*reinterpret_cast<const uint64_t *>(s_y + j)), "l"(qyp));
                asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(dz) : "l"(*reinterpret_cast<const uint64_t *>(s_z + j)), "l"(qzp));
                asm("mul.rn.f32x2 %0, %1, %1;" : "=l"(acc) : "l"(dx));
                asm("fma.rn.f32x2 %0, %1, %1, %2;" : "=l"(acc) : "l"(dy), "l"(acc));
                asm("fma.rn.f32x2 %0, %1, %1, %2;" : "=l"(acc) : "l"(dz), "l"(acc));
                asm("mov.b64 {%0, %1}, %2;" : "=f"(d[2 * i]), "=f"(d[2 * i + 1]) : "l"(acc));
                asm("min.f32 %0, %0, %1, %2;" : "+f"(lmin) : "f"(d[2 * i]), "f"(d[2 * i + 1]));
            }
        }
        // threshold: k rounds of "take the smallest remaining lane minimum" (d >= 0, so its bit image is ordered).
        // With k close to 32 the k-th smallest of 32 lane minima is a weak bound (k = 30 on 512 candidates: ~100
        // survivors, the slow selection below); there every lane offers its TWO smallest distances instead, and the
        // k-th smallest of those 64 real candidates is close to the true k-th distance (~35 survivors).
        uint32_t v = __float_as_uint(lmin), v2 = 0xffffffffu, T = 0;
        if (k > KNNF_TWOMIN_K) {                                // warp-uniform
            float m1 = INFINITY, m2 = INFINITY;
#pragma unroll
            for (int i = 0; i < KPL; ++i) {
                m2 = fminf(m2, fmaxf(m1, d[i]));
                m1 = fminf(m1, d[i]);
            }
            v2 = __float_as_uint(m2);
        }
        for (int r = 0; r < k; ++r) {
            T = __reduce_min_sync(0xffffffffu, v);
            if (v == T) { v = v2; v2 = 0xffffffffu; }
        }
        // count + compact everything with d <= T
        int cnt = 0;
#pragma unroll
        for (int i = 0; i < KPL; ++i) cnt += (__float_as_uint(d[i]) <= T) ? 1 : 0;
        int off = cnt;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, off, s);
            if (lane >= s) off += o;
        }
        const int total = __shfl_sync(0xffffffffu, off, 31);
        off -= cnt;
        uint64_t *cand = s_cand[warp];
        uint64_t c0 = KMAX, c1 = KMAX, c2 = KMAX, c3 = KMAX;
        const bool fits = total <= KNNF_CAP;                    // warp-uniform
        __syncwarp();
        if (fits) {
#pragma unroll
            for (int i = 0; i < KPL; ++i)
                if (__float_as_uint(d[i]) <= T)
                    cand[off++] = ((uint64_t)__float_as_uint(d[i]) << 32) | (uint32_t)(2 * ((i >> 1) * 32 + lane) + (i & 1));
            __syncwarp();
            c0 = lane < total ? cand[lane] : KMAX;
            c1 = lane + 32 < total ? cand[lane + 32] : KMAX;
            c2 = lane + 64 < total ? cand[lane + 64] : KMAX;
            c3 = lane + 96 < total ? cand[lane + 96] : KMAX;
        }
        (void)lt_mask;
        const size_t o0 = gy * k;
        if (total <= 64) {                                      // warp-uniform; the usual case (~1.5 k survivors)
            // rank by counting: keys are unique (index in the low word), so #(keys smaller than mine) is my place in
            // the ascending (d2, index) order -- one broadcast shared-memory read per survivor instead of k ordered
            // extractions with two 64-bit warp reductions each
            int r0 = 0, r1 = 0;
#pragma unroll 4
            for (int j = 0; j < total; ++j) {
                const uint64_t kj = cand[j];
                r0 += kj < c0 ? 1 : 0;
                r1 += kj < c1 ? 1 : 0;
            }
            if (lane < total && r0 < k) {
                const uint32_t ci = (uint32_t)c0;
                if (row_out) row_out[o0 + r0] = (int64_t)gy;
                if (col_out) col_out[o0 + r0] = (int64_t)((size_t)b * Nx + ci);
                if (col_local) col_local[o0 + r0] = (int32_t)ci;
                if (nbr4) nbr4[o0 + r0] = make_float4(s_x[ci] - qx, s_y[ci] - qy, s_z[ci] - qz, __int_as_float((int)ci));
            }
            if (lane + 32 < total && r1 < k) {
                const uint32_t ci = (uint32_t)c1;
                if (row_out) row_out[o0 + r1] = (int64_t)gy;
                if (col_out) col_out[o0 + r1] = (int64_t)((size_t)b * Nx + ci);
                if (col_local) col_local[o0 + r1] = (int32_t)ci;
                if (nbr4) nbr4[o0 + r1] = make_float4(s_x[ci] - qx, s_y[ci] - qy, s_z[ci] - qz, __int_as_float((int)ci));
            }
            __syncwarp();
            continue;
        }
        for (int r = 0; r < k; ++r) {
            uint64_t lm;
            if (fits) {
                lm = c0 < c1 ? c0 : c1;
                const uint64_t t2 = c2 < c3 ? c2 : c3;
                lm = lm < t2 ? lm : t2;
            } else {                                            // rare: rescan the full register set
                lm = KMAX;
#pragma unroll
                for (int i = 0; i < KPL; ++i) {
                    const uint64_t key = ((uint64_t)__float_as_uint(d[i]) << 32) | (uint32_t)(2 * ((i >> 1) * 32 + lane) + (i & 1));
                    lm = (d[i] != INFINITY && key < lm) ? key : lm;
                }
            }
            const uint64_t win = warp_min_u64(lm);
            if (lm == win) {                                    // exactly one lane: indices are unique
                if (fits) {
                    c0 = c0 == win ? KMAX : c0; c1 = c1 == win ? KMAX : c1;
                    c2 = c2 == win ? KMAX : c2; c3 = c3 == win ? KMAX : c3;
                } else {
#pragma unroll
                    for (int i = 0; i < KPL; ++i)
                        if ((uint32_t)win == (uint32_t)(2 * ((i >> 1) * 32 + lane) + (i & 1))) d[i] = INFINITY;
                }
            }
            if (lane == 0) {
                const uint32_t ci = (uint32_t)win;
                if (row_out) row_out[o0 + r] = (int64_t)gy;
                if (col_out) col_out[o0 + r] = (int64_t)((size_t)b * Nx + ci);
                if (col_local) col_local[o0 + r] = (int32_t)ci;
                if (nbr4) nbr4[o0 + r] = make_float4(s_x[ci] - qx, s_y[ci] - qy, s_z[ci] - qz, __int_as_float((int)ci));
            }
        }
        __syncwarp();
    }
}

int knn_launch(const float *x, const float *y, int B, int Nx, int Ny, int k, int64_t *row, int64_t *col,
               int32_t *col_local, float4 *nbr4, cudaStream_t st) {
    const dim3 fgrid(div_up(Ny, KNNF_WARPS * KNNF_QPW), B);
    if (k <= 32 && Nx <= 512) {
        knn_fast_kernel<16><<<fgrid, KNNF_WARPS * 32, 0, st>>>(x, y, Nx, Ny, k, row, col, col_local, nbr4);
        DCLR_LAUNCH_CHECK("knn_fast_kernel");
    } else if (k <= 32 && Nx <= 1024) {
        knn_fast_kernel<32><<<fgrid, KNNF_WARPS * 32, 0, st>>>(x, y, Nx, Ny, k, row, col, col_local, nbr4);
        DCLR_LAUNCH_CHECK("knn_fast_kernel");
    } else {
        knn_kernel<<<dim3(div_up(Ny, KNN_WARPS * KNN_QPW), B), KNN_WARPS * 32, 0, st>>>(x, y, Nx, Ny, k, row, col, col_local, nbr4);
        DCLR_LAUNCH_CHECK("knn_kernel");
    }
    return DCLR_OK;
}

}  // namespace dclr

// ------------------------------------------------------------------ C ABI
extern "C" int dclr_gather(const float *feat, const int32_t *idx, int B, int C, int N, int M, float *out,
                           void *stream) {
    using namespace dclr;
    DCLR_CHECK_ARG(feat && idx && out, "dclr_gather: null pointer");
    DCLR_CHECK_ARG(B >= 0 && C >= 1 && N >= 1 && M >= 0 && C <= 65535 && B <= 65535, "dclr_gather: bad shape");
    if (B == 0 || M == 0) return DCLR_OK;
    gather_kernel<<<dim3(div_up(M, 256), C, B), 256, 0, (cudaStream_t)stream>>>(feat, idx, C, N, M, out);
    DCLR_LAUNCH_CHECK("gather_kernel");
    return DCLR_OK;
}

extern "C" int dclr_group(const float *feat, const int32_t *idx, int B, int C, int N, int M, int nsample, float *out,
                          void *stream) {
    using namespace dclr;
    DCLR_CHECK_ARG(feat && idx && out, "dclr_group: null pointer");
    DCLR_CHECK_ARG(B >= 0 && C >= 1 && N >= 1 && M >= 0 && nsample >= 0 && C <= 65535 && B <= 65535,
                   "dclr_group: bad shape");
    if (B == 0 || M == 0 || nsample == 0) return DCLR_OK;
    size_t total = (size_t)M * nsample;
    group_kernel<<<dim3((unsigned)((total + 255) / 256), C, B), 256, 0, (cudaStream_t)stream>>>(feat, idx, C, N, M,
                                                                                                nsample, out);
    DCLR_LAUNCH_CHECK("group_kernel");
    return DCLR_OK;
}

extern "C" int dclr_gather_grad(const float *grad_out, const int32_t *idx, int B, int C, int N, int M,
                                float *grad_points, void *stream) {
    using namespace dclr;
    DCLR_CHECK_ARG(grad_out && idx && grad_points, "dclr_gather_grad: null pointer");
    DCLR_CHECK_ARG(B >= 0 && C >= 1 && N >= 1 && M >= 0 && C <= 65535 && B <= 65535, "dclr_gather_grad: bad shape");
    if (B == 0 || M == 0) return DCLR_OK;
    gather_grad_kernel<<<dim3(div_up(M, 256), C, B), 256, 0, (cudaStream_t)stream>>>(grad_out, idx, C, N, M, grad_points);
    DCLR_LAUNCH_CHECK("gather_grad_kernel");
    return DCLR_OK;
}

extern "C" int dclr_group_grad(const float *grad_out, const int32_t *idx, int B, int C, int N, int M, int nsample,
                               float *grad_points, void *stream) {
    using namespace dclr;
    DCLR_CHECK_ARG(grad_out && idx && grad_points, "dclr_group_grad: null pointer");
    DCLR_CHECK_ARG(B >= 0 && C >= 1 && N >= 1 && M >= 0 && nsample >= 0 && C <= 65535 && B <= 65535,
                   "dclr_group_grad: bad shape");
    if (B == 0 || M == 0 || nsample == 0) return DCLR_OK;
    size_t total = (size_t)M * nsample;
    group_grad_kernel<<<dim3((unsigned)((total + 255) / 256), C, B), 256, 0, (cudaStream_t)stream>>>(
        grad_out, idx, C, N, M, nsample, grad_points);
    DCLR_LAUNCH_CHECK("group_grad_kernel");
    return DCLR_OK;
}

extern "C" int dclr_ball_query(const float *new_xyz, const float *pts, int pt_stride, int B, int N, int M,
                               float radius, int nsample, int32_t *idx_out, int32_t *cnt_out, int pad, void *stream) {
    using namespace dclr;
    DCLR_CHECK_ARG(new_xyz && pts && idx_out, "dclr_ball_query: null pointer");
    DCLR_CHECK_ARG(pt_stride >= 3 && B >= 0 && N >= 1 && M >= 0 && nsample >= 1 && B <= 65535,
                   "dclr_ball_query: bad shape");
    if (B == 0 || M == 0) return DCLR_OK;
    int32_t *idx[1] = {idx_out};
    int32_t *cnt[1] = {cnt_out};
    cudaStream_t st = (cudaStream_t)stream;
    if (N >= BG_MIN_N) {                                   // large clouds: uniform-grid candidate pruning (ballgrid.cu)
        void *ws = nullptr;
        { int rc_ = scratch_alloc(&ws, ball_grid_workspace_bytes(B, N), st); if (rc_ != DCLR_OK) return rc_; }
        int used = 0;
        int rc = ball_grid_launch(new_xyz, pts, pt_stride, B, N, M, 1, &radius, &nsample, idx, cnt, pad, ws, st, &used);
        scratch_free(ws, st);
        if (rc != DCLR_OK || used) return rc;
    }
    return ball_query_launch(new_xyz, pts, pt_stride, B, N, M, 1, &radius, &nsample, idx, cnt, pad,
                             (cudaStream_t)stream);
}

extern "C" int dclr_knn(const float *x, const float *y, int B, int Nx, int Ny, int k, int64_t *row_out,
                        int64_t *col_out, void *stream) {
    using namespace dclr;
    DCLR_CHECK_ARG(x && y && row_out && col_out, "dclr_knn: null pointer");
    DCLR_CHECK_ARG(B >= 0 && Nx >= 1 && Ny >= 0 && B <= 65535, "dclr_knn: bad shape");
    DCLR_CHECK_ARG(k >= 1 && k <= Nx && k <= KNN_MAXK, "dclr_knn: need 1 <= k <= min(Nx,%d), got k=%d Nx=%d", KNN_MAXK,
                   k, Nx);
    if (B == 0 || Ny == 0) return DCLR_OK;
    return knn_launch(x, y, B, Nx, Ny, k, row_out, col_out, nullptr, nullptr, (cudaStream_t)stream);
}
