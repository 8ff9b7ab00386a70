// 16-bit tcgen05/TMEM path for the two dense stacks of the forward:
//   flow_tc_kernel : kNN-gathered rows [f_tmpl | f_src | dp] -> 3 x (linear + bias + ReLU) -> radius mask -> max over k
//                    (reference deepclr/models/deepclr.py:201-231; rows are built in shared memory, never in HBM)
//   head_tc_kernel : [motion feat | xyz] -> 5 x (linear + bias + ReLU) -> max over the n points   (deepclr.py:284-287)
// Operands are bf16 (or fp16) in the canonical no-swizzle K-major UMMA layout, accumulation is fp32 in TMEM.
//
// Shared-memory operand layout (both activations and weights): an [R rows x K] tile is stored as K/8 "chunks", each
// chunk holding 8 consecutive K-elements (16 bytes) of every row, rows consecutive:
//     byte offset(r, k) = (k / 8) * (R * 16) + r * 16 + (k % 8) * 2
// i.e. 8x8 core matrices of 128 contiguous bytes, SBO (next 8-row group) = 128 B, LBO (next K chunk) = R * 16 B.
// An epilogue thread that owns row r therefore writes one 16-byte vector per chunk and a warp writes 512 contiguous
// bytes: conflict free.  Weights are pre-packed into exactly this image at model creation so that a weight tile is
// one contiguous cp.async.bulk (TMA engine) transfer.
//
// Hidden layers compute D[rows, out] = X[rows, K] * W[out, K]^T (A = activations, thread <-> row in the epilogue).
// The LAST layer of each stack is computed transposed, D^T[out, rows] = W * X^T (A = weights): TMEM lanes are output
// channels and TMEM columns are rows, so the max over the k neighbours / over the points is a per-thread reduction
// over registers -- no shuffles, no shared memory.  Because both operands use the same K-major tile image, swapping
// the roles costs nothing.
#include "common.cuh"
#include "kernels.cuh"
#include "mlp_tc.cuh"
#include "ptx.cuh"
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <new>
#include <vector>
#include <algorithm>
#include <type_traits>

namespace dclr {

constexpr int TC_ROWS = 128;          // rows (UMMA M or N) of every activation tile
constexpr int TC_LBO = TC_ROWS * 16;  // bytes between K chunks of a 128-row tile

// ---------------------------------------------------------------- descriptors
// Instruction descriptor for tcgen05.mma.kind::f16 (cute::UMMA::InstrDescriptor): c_format[4,6)=1 (f32),
// a_format[7,10) / b_format[10,13) = 0 f16 | 1 bf16, a_major bit15 = b_major bit16 = 0 (K-major), n_dim[17,23) = N>>3,
// m_dim[24,29) = M>>4.
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, int is_bf16) {
    return (1u << 4) | ((uint32_t)is_bf16 << 7) | ((uint32_t)is_bf16 << 10) | ((uint32_t)(N >> 3) << 17) |
           ((uint32_t)(M >> 4) << 24);
}
// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor), SWIZZLE_NONE, K-major: start>>4 in [0,14),
// LBO>>4 in [16,30), SBO>>4 in [32,46), version=1 in [46,48).
__device__ __forceinline__ uint64_t make_sdesc(uint32_t smem_addr, uint32_t lbo_bytes) {
    return (uint64_t)((smem_addr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
           ((uint64_t)(128u >> 4) << 32) | (1ull << 46);
}

// Descriptors in the issue loops: constant high word, low word advanced by an ADD per K step (the issuing warp's
// integer work per MMA is serial latency in front of the tensor pipe).
constexpr uint32_t SDESC_HI = (128u >> 4) | (1u << 14);                 // SBO = 128 B, version 1 (bits 32..47)
constexpr uint32_t SDESC_KSTEP = (2u * TC_LBO) >> 4;                    // two 16-byte chunk columns = 16 K elements
__device__ __forceinline__ uint32_t sdesc_lo(uint32_t smem_addr) {     // start address | LBO = TC_LBO
    return ((smem_addr >> 4) & 0x3fffu) | ((uint32_t)(TC_LBO >> 4) << 16);
}
__device__ __forceinline__ uint64_t sdesc(uint32_t lo) { return ((uint64_t)SDESC_HI << 32) | lo; }

template <typename T16> struct Cvt;
template <> struct Cvt<__nv_bfloat16> {
    static constexpr int is_bf16 = 1;
    static __device__ __forceinline__ uint32_t pack(float a, float b) {
        __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t *>(&v);
    }
    static __device__ __forceinline__ uint32_t pack_relu(uint32_t a, uint32_t b) {     // max(.,0) then round
        uint32_t r;
        asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(__uint_as_float(b)), "f"(__uint_as_float(a)));
        return r;
    }
    static __device__ __forceinline__ uint16_t one(float a) {
        __nv_bfloat16 v = __float2bfloat16_rn(a);
        return *reinterpret_cast<uint16_t *>(&v);
    }
    static __device__ __forceinline__ float back(uint16_t h) { return __uint_as_float((uint32_t)h << 16); }
};
template <> struct Cvt<__half> {
    static constexpr int is_bf16 = 0;
    static __device__ __forceinline__ uint32_t pack(float a, float b) {
        // saturate instead of overflowing to inf
        a = fminf(fmaxf(a, -65504.f), 65504.f);
        b = fminf(fmaxf(b, -65504.f), 65504.f);
        __half2 v = __floats2half2_rn(a, b);
        return *reinterpret_cast<uint32_t *>(&v);
    }
    static __device__ __forceinline__ uint32_t pack_relu(uint32_t a, uint32_t b) {     // max(.,0), round, saturate
        uint32_t r;
        asm("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(__uint_as_float(b)), "f"(__uint_as_float(a)));
        return r;
    }
    static __device__ __forceinline__ uint16_t one(float a) {
        a = fminf(fmaxf(a, -65504.f), 65504.f);
        __half v = __float2half_rn(a);
        return *reinterpret_cast<uint16_t *>(&v);
    }
    static __device__ __forceinline__ float back(uint16_t h) { return __half2float(*reinterpret_cast<__half *>(&h)); }
};

__device__ __forceinline__ void ld_shared_v4(uint32_t addr, uint32_t &a, uint32_t &b, uint32_t &c, uint32_t &d) {
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr) : "memory");
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// ---------------------------------------------------------------- weight packing (model creation time)
// dst image of rows [row0, row0+R) x padded K [k0, k0+Kt) of W[out,in] (fp32, row-major) in the chunk layout above.
// nfeat == 0: plain copy, zero padded.  nfeat > 0: the reference puts the 3 coordinates first (deepclr.py:208-210,228);
// here the nfeat feature columns come first (16-byte aligned chunks) and are followed by NINE coordinate columns
//     [w_hi(x,y,z) | w_lo(x,y,z) | w_hi(x,y,z)]   against activations   [c_hi(x,y,z) | c_hi(x,y,z) | c_lo(x,y,z)]
// (hi = value rounded to 16 bits, lo = rounded remainder), i.e. the coordinate contribution w*c is evaluated as
// w_hi*c_hi + w_lo*c_hi + w_hi*c_lo with ~16 significant bits: metre-scale coordinates (up to 80 m on LiDAR scans)
// would otherwise lose centimetres to the 8-bit bf16 mantissa.  The columns live in what is padding anyway.
// bias (nullable) folds the layer bias into the GEMM: columns kbias / kbias + 1 hold b_hi / b_lo and meet two columns
// of ONES in the activations, so the accumulator already contains x.W + b (to ~16 significant bits of b) and the
// hidden-layer epilogue is a single cvt.rn.relu per value pair.
template <typename T16>
__global__ void pack_tile_kernel(const float *__restrict__ W, const float *__restrict__ bias, int kbias, int out, int in,
                                 int nfeat, int row0, int R, int k0, int Kt, uint16_t *__restrict__ dst) {
    int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= R * Kt) return;
    int kk = e % Kt, r = e / Kt;
    int kp = k0 + kk, row = row0 + r;
    uint16_t q = 0;
    if (row < out && bias != nullptr && (kp == kbias || kp == kbias + 1)) {
        const float bv = bias[row];
        const uint16_t hi = Cvt<T16>::one(bv);
        q = kp == kbias ? hi : Cvt<T16>::one(bv - Cvt<T16>::back(hi));
    } else if (row < out) {
        if (nfeat == 0) {
            if (kp < in) q = Cvt<T16>::one(W[(size_t)row * in + kp]);
        } else if (kp < nfeat) {
            q = Cvt<T16>::one(W[(size_t)row * in + kp + 3]);
        } else if (kp < nfeat + 9) {
            const int j = kp - nfeat;
            const float w = W[(size_t)row * in + j % 3];
            const uint16_t hi = Cvt<T16>::one(w);
            q = (j / 3 == 1) ? Cvt<T16>::one(w - Cvt<T16>::back(hi)) : hi;
        }
    }
    dst[(size_t)(kk / 8) * (R * 8) + r * 8 + (kk % 8)] = q;
}

// the two 16-byte chunks holding a coordinate triple as [hi hi lo] (see pack_tile_kernel)
template <typename T16>
__device__ __forceinline__ void store_coord_chunks(uint32_t addr0, uint32_t addr1, float x, float y, float z) {
    const uint16_t xh = Cvt<T16>::one(x), yh = Cvt<T16>::one(y), zh = Cvt<T16>::one(z);
    const uint16_t xl = Cvt<T16>::one(x - Cvt<T16>::back(xh)), yl = Cvt<T16>::one(y - Cvt<T16>::back(yh)),
                   zl = Cvt<T16>::one(z - Cvt<T16>::back(zh));
    st_shared_v4(addr0, xh | ((uint32_t)yh << 16), zh | ((uint32_t)xh << 16), yh | ((uint32_t)zh << 16),
                 xl | ((uint32_t)yl << 16));
    const uint32_t one = Cvt<T16>::one(1.0f);
    st_shared_v4(addr1, zl | (one << 16), one, 0u, 0u);          // [z_lo, 1, 1, 0 ...]: the ones meet [b_hi, b_lo]
}

// ---------------------------------------------------------------- shared epilogue pieces
// Hidden-layer epilogue of one thread (= one tile row): TMEM accumulators [ncols] (bias already inside, see
// pack_tile_kernel) -> ReLU + round to 16 bits in ONE instruction per pair -> the row's 16-byte vectors of the next
// layer's A/B operand tile.  taddr already contains the warp's lane offset.
template <typename T16>
__device__ __forceinline__ void epilogue_hidden(uint32_t taddr, int c_begin, int c_end, uint32_t act_row_addr) {
    for (int c0 = c_begin; c0 < c_end; c0 += 32) {
        uint32_t v[32];
        ptx::tmem_ld32(taddr + c0, v);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 4; ++q)
            st_shared_v4(act_row_addr + (uint32_t)((c0 >> 3) + q) * TC_LBO,
                         Cvt<T16>::pack_relu(v[q * 8 + 0], v[q * 8 + 1]), Cvt<T16>::pack_relu(v[q * 8 + 2], v[q * 8 + 3]),
                         Cvt<T16>::pack_relu(v[q * 8 + 4], v[q * 8 + 5]), Cvt<T16>::pack_relu(v[q * 8 + 6], v[q * 8 + 7]));
    }
}
// the two chunks after a layer's `width` output columns: [1, 1, 0 ...] against the next layer's [b_hi, b_lo, 0 ...]
template <typename T16>
__device__ __forceinline__ void store_ones_chunks(uint32_t act_row_addr, int width) {
    const uint32_t one = Cvt<T16>::one(1.0f);
    st_shared_v4(act_row_addr + (uint32_t)(width >> 3) * TC_LBO, one | (one << 16), 0u, 0u, 0u);
    st_shared_v4(act_row_addr + (uint32_t)((width >> 3) + 1) * TC_LBO, 0u, 0u, 0u, 0u);
}

#ifndef DCLR_FLOW_EPI3_COLS
#define DCLR_FLOW_EPI3_COLS 64
#endif
constexpr int FLOW_EPI3_COLS = DCLR_FLOW_EPI3_COLS;   // TMEM columns in flight per wait in the last flow epilogue (64 | 32)
// Last flow layer (transposed): this thread's TMEM lane is one output channel, the 128 columns are the tile rows
// (groups of k neighbours).  Radius-masked max per group, + bias, ReLU, store.  KC > 0: k known at compile time, the
// group bookkeeping folds away and the per-group maxima become independent chains.
template <int KC, typename T16>
__device__ __forceinline__ void flow_group_max(uint32_t taddr, const uint32_t (&mk)[4], int k_rt, int g, int nvalid,
                                               float bias, uint16_t *__restrict__ out, float *__restrict__ out32,
                                               int ld) {
    if constexpr (KC > 0) {
        // k known at compile time: column -> group is a constant, so the group maxima are G independent register
        // chains (no counters, no branches); the stores follow as one short predicated sequence
        constexpr int G = 128 / KC;
        float gm[G];
#pragma unroll
        for (int q = 0; q < G; ++q) gm[q] = -INFINITY;
        const bool any_masked = (mk[0] | mk[1] | mk[2] | mk[3]) != 0u;   // same for every thread of the slot
#pragma unroll
        for (int c0 = 0; c0 < 128; c0 += FLOW_EPI3_COLS) {
            uint32_t v0[32], v1[32];
            ptx::tmem_ld32(taddr + c0, v0);
            if constexpr (FLOW_EPI3_COLS == 64) ptx::tmem_ld32(taddr + c0 + 32, v1);
            ptx::tmem_ld_wait();
            if (!any_masked) {
#pragma unroll
                for (int i = 0; i < FLOW_EPI3_COLS; ++i) {
                    const int col = c0 + i;
                    if (col / KC < G) gm[col / KC] = fmaxf(gm[col / KC], __uint_as_float(i < 32 ? v0[i & 31] : v1[i & 31]));
                }
            } else {
#pragma unroll
                for (int i = 0; i < FLOW_EPI3_COLS; ++i) {
                    const int col = c0 + i;
                    const bool m = (mk[col >> 5] >> (col & 31)) & 1u;
                    const float x = m ? -INFINITY : __uint_as_float(i < 32 ? v0[i & 31] : v1[i & 31]);
                    if (col / KC < G) gm[col / KC] = fmaxf(gm[col / KC], x);
                }
            }
        }
        const int ng = min(min(g, nvalid), G);
#pragma unroll
        for (int q = 0; q < G; ++q) {
            if (q < ng) {
                const float val = fmaxf(gm[q] + bias, 0.f);
                out[(size_t)q * ld] = Cvt<T16>::one(val);
                if (out32) out32[(size_t)q * ld] = val;
            }
        }
        return;
    }
    const int k = k_rt;
    float cur = -INFINITY;
    int cnt = 0, og = 0;
    auto scan = [&](const uint32_t (&v)[32], uint32_t mword) {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const float x = ((mword >> i) & 1u) ? -INFINITY : __uint_as_float(v[i]);
            cur = fmaxf(cur, x);
            if (++cnt == k) {
                if (og < g && og < nvalid) {
                    const float val = fmaxf(cur + bias, 0.f);
                    out[(size_t)og * ld] = Cvt<T16>::one(val);
                    if (out32) out32[(size_t)og * ld] = val;
                }
                ++og; cnt = 0; cur = -INFINITY;
            }
        }
    };
#pragma unroll 1
    for (int c0 = 0; c0 < 128; c0 += 32) {
        uint32_t v[32];
        ptx::tmem_ld32(taddr + c0, v);
        ptx::tmem_ld_wait();
        scan(v, mk[c0 >> 5]);
    }
}

// ---------------------------------------------------------------- flow embedding
#ifdef DCLR_TRACE
// tools/head_trace.py: SM-clock timestamps of CTA 0's roles (build -DDCLR_TRACE only)
__device__ long long g_trace[1024];
__device__ long long g_flow_trace[1024];
#define FLTRACE(slot) do { if (blockIdx.x == 0 && lane == 0 && q == 0 && half == 0 && s == 0 && ft >= 0 && ft < 8) g_flow_trace[ft * 16 + (slot)] = clock64(); } while (0)
#define TRACE(slot) do { if (blockIdx.x == 0 && (slot) < 1024) g_trace[(slot)] = clock64(); } while (0)
#else
#define FLTRACE(slot) do { } while (0)
#define TRACE(slot) do { } while (0)
#endif
struct FlowArgs {
    const uint16_t *feat;   // [2B, n, F] point-major set-abstraction features, 16-bit operand format, templates first
    const float4 *nbr;      // [B, n, k] per row: {source - template displacement, source index (local to the pair)} of
                            // the k nearest source samples, written by the kNN kernel (pointops.cu)
    const uint16_t *w;      // packed images: W1 [128 x K0p] | W2 [128 x 128] | W3 [H3 x 128] (128-row blocks)
    const float *bias;      // b1[128] b2[128] b3[H3]
    uint16_t *out;          // [B*n, H3] motion features, 16-bit (coordinates are kept in xyz)
    float *out32;           // same in fp32 for the API tap, or null
    int B, n, k, F, K0p, H3;
    float radius;
    int groups_per_tile, tiles_per_pair, num_tiles;
};

constexpr int FLOW_SLOTS = 2;
constexpr int FLOW_K2P = 128 + 16;     // layer-2 K: 128 hidden + [b_hi, b_lo, 0 ...] against the ones chunks
constexpr int FLOW_WPS = 8;           // build/epilogue warps per slot: 4 TMEM lane quarters x 2 column halves
constexpr int FLOW_MMA_WARP = FLOW_SLOTS * FLOW_WPS;
#ifndef DCLR_FLOW_STAGE
#define DCLR_FLOW_STAGE 0
#endif
#ifndef DCLR_FLOW_PROD
#define DCLR_FLOW_PROD 1
#endif
constexpr int FLOW_STAGE_BYTES = DCLR_FLOW_STAGE;   // per gather warp: staging area for a tile's distinct template rows
constexpr int FLOW_PROD = DCLR_FLOW_PROD;           // gather (cp.async) warps per slot; 19 warps = at most 5 per SM
                                                    // sub-partition = 96 registers per thread
constexpr int FLOW_THREADS = (FLOW_MMA_WARP + 1 + FLOW_SLOTS * FLOW_PROD) * 32;

template <typename T16>
__global__ void __launch_bounds__(FLOW_THREADS, 1)
flow_tc_kernel(FlowArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int K0p = a.K0p, H3 = a.H3;
    const uint32_t w1_bytes = 128u * K0p * 2, w2_bytes = 128u * FLOW_K2P * 2, w3_bytes = (uint32_t)H3 * 128 * 2;
    const uint32_t act_bytes = 128u * K0p * 2;
    uint8_t *s_w1 = smem, *s_w2 = s_w1 + w1_bytes, *s_w3 = s_w2 + w2_bytes;
    uint8_t *s_act = s_w3 + w3_bytes;                                    // FLOW_SLOTS x act_bytes
    float *s_bias = reinterpret_cast<float *>(s_act + FLOW_SLOTS * act_bytes);   // 256 + H3 floats
    float4 *s_nbr = reinterpret_cast<float4 *>(s_bias + 256 + H3);               // [FLOW_SLOTS][128] row records
    uint8_t *s_stage = reinterpret_cast<uint8_t *>(s_nbr + FLOW_SLOTS * 128);    // per gather warp: template rows of a tile
    uint32_t *s_maskw = reinterpret_cast<uint32_t *>(s_stage + FLOW_SLOTS * FLOW_PROD * FLOW_STAGE_BYTES);   // [2 parities][FLOW_SLOTS][4]
    uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_maskw + 2 * FLOW_SLOTS * 4);    // wbar, ready, done, idx, free, gath
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(s_bar + 1 + 5 * FLOW_SLOTS);
    uint64_t *wbar = s_bar, *ready = s_bar + 1, *done = s_bar + 1 + FLOW_SLOTS, *idx_ready = s_bar + 1 + 2 * FLOW_SLOTS;
    uint64_t *act_free = s_bar + 1 + 3 * FLOW_SLOTS, *gath = s_bar + 1 + 4 * FLOW_SLOTS;

    if (tid == 0) {
        ptx::mbar_init(wbar, 1);
        for (int s = 0; s < FLOW_SLOTS; ++s) {
            ptx::mbar_init(&ready[s], FLOW_WPS * 32);
            ptx::mbar_init(&done[s], 1);
            ptx::mbar_init(&idx_ready[s], 32);
            ptx::mbar_init(&act_free[s], 1);
            ptx::mbar_init(&gath[s], FLOW_PROD * 32);
        }
        ptx::fence_mbar_init();
    }
    if (warp == FLOW_MMA_WARP) ptx::tmem_alloc(s_tmem, 512);
    for (int e = tid; e < 256 + H3; e += FLOW_THREADS) s_bias[e] = __ldg(a.bias + e);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *s_tmem;
    const int tiles_per_iter = gridDim.x * FLOW_SLOTS;
    constexpr uint32_t IDESC = make_idesc(128, 128, Cvt<T16>::is_bf16);

    if (warp == FLOW_MMA_WARP) {
        // ===================== weight load (once) + MMA issue =====================
        // The row records of a slot's NEXT tile ({displacement, source index}, 16 bytes per row) travel global -> shared
        // by cp.async from this (mostly idle) warp: no other warp of the kernel loads coordinates or indices from
        // global memory.  The 128 records of a tile are contiguous: nbr[(b n + p0) k + r], r < g k.  Always one
        // arrival per lane, copy or not.
        auto prefetch_idx = [&](const int tile, const int s) {
            if (tile < a.num_tiles) {
                const int b = tile / a.tiles_per_pair;
                const int p0 = (tile - b * a.tiles_per_pair) * a.groups_per_tile;
                const int rows = min(a.groups_per_tile, a.n - p0) * a.k;
                const float4 *src = a.nbr + ((size_t)b * a.n + p0) * a.k;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = i * 32 + lane;
                    if (r < rows) ptx::cp_async16_cg(ptx::smem_u32(s_nbr + s * 128 + r), src + r, 16);
                }
            }
            ptx::cp_async_mbar_arrive_noinc(&idx_ready[s]);
        };
#pragma unroll
        for (int s = 0; s < FLOW_SLOTS; ++s) prefetch_idx(blockIdx.x * FLOW_SLOTS + s, s);
        if (lane == 0) {
            ptx::mbar_expect_tx(wbar, w1_bytes + w2_bytes + w3_bytes);
            ptx::bulk_g2s(s_w1, a.w, w1_bytes, wbar);
            ptx::bulk_g2s(s_w2, reinterpret_cast<const uint8_t *>(a.w) + w1_bytes, w2_bytes, wbar);
            ptx::bulk_g2s(s_w3, reinterpret_cast<const uint8_t *>(a.w) + w1_bytes + w2_bytes, w3_bytes, wbar);
            ptx::mbar_wait(wbar, 0);
        }
        __syncwarp();
        const uint32_t w1_lo = sdesc_lo(ptx::smem_u32(s_w1)), w2_lo = sdesc_lo(ptx::smem_u32(s_w2)),
                       w3_lo = sdesc_lo(ptx::smem_u32(s_w3));
        uint32_t act_lo[FLOW_SLOTS];
#pragma unroll
        for (int s = 0; s < FLOW_SLOTS; ++s) act_lo[s] = sdesc_lo(ptx::smem_u32(s_act + s * act_bytes));
        // The whole warp runs the issue loop (converged); one elected lane issues each MMA / commit (ptx.cuh).
        // Slots are served in the order they become ready, not round-robin: a fixed s0, s1, s0, ... order locks the two
        // slots into the same phase (each waited ~1150 extra cycles per layer for the other's MMAs, both gathered and
        // both reduced at the same time, tensor pipe idle meanwhile); served on demand they drift into anti-phase.
        uint32_t rphase[FLOW_SLOTS], layer[FLOW_SLOTS];
        int tile[FLOW_SLOTS];
        bool live[FLOW_SLOTS];
        bool any_live = false;
#pragma unroll
        for (int s = 0; s < FLOW_SLOTS; ++s) {
            rphase[s] = 0; layer[s] = 0;
            tile[s] = blockIdx.x * FLOW_SLOTS + s;
            live[s] = tile[s] < a.num_tiles;
            any_live |= live[s];
        }
        while (any_live) {
#pragma unroll
            for (int s = 0; s < FLOW_SLOTS; ++s) {
                if (!live[s]) continue;
                if (!__all_sync(0xffffffffu, ptx::mbar_try_wait(&ready[s], rphase[s] & 1))) continue;   // warp-uniform
                ptx::tc_fence_after();
                const uint32_t dcol = tmem_base + s * 256;
                const uint32_t a_lo = act_lo[s];
                if (layer[s] == 0) {
                    for (int ks = 0; ks < K0p / 16; ++ks)
                        ptx::mma_f16_ss_warp(dcol, sdesc(a_lo + ks * SDESC_KSTEP), sdesc(w1_lo + ks * SDESC_KSTEP), IDESC,
                                             ks > 0);
                } else if (layer[s] == 1) {
#pragma unroll
                    for (int ks = 0; ks < FLOW_K2P / 16; ++ks)
                        ptx::mma_f16_ss_warp(dcol + 128, sdesc(a_lo + ks * SDESC_KSTEP), sdesc(w2_lo + ks * SDESC_KSTEP),
                                             IDESC, ks > 0);
                } else {
                    for (int mb = 0; mb < H3 / 128; ++mb) {
                        const uint32_t w_lo = w3_lo + mb * ((128u * 128 * 2) >> 4);
#pragma unroll
                        for (int ks = 0; ks < 8; ++ks)      // transposed: A = weights, B = activations
                            ptx::mma_f16_ss_warp(dcol + mb * 128, sdesc(w_lo + ks * SDESC_KSTEP),
                                                 sdesc(a_lo + ks * SDESC_KSTEP), IDESC, ks > 0);
                    }
                }
                ptx::mma_commit_warp(&done[s]);
                ++rphase[s];
                if (layer[s] == 0) {
                    // every thread of slot s has read its index (they arrived on ready[s]): refill for the next tile
                    prefetch_idx(tile[s] + tiles_per_iter, s);
                    layer[s] = 1;
                } else if (layer[s] == 1) {
                    layer[s] = 2;
                } else {
                    ptx::mma_commit_warp(&act_free[s]);                // operand buffer free: the gather warp goes on
                    layer[s] = 0;
                    tile[s] += tiles_per_iter;
                    live[s] = tile[s] < a.num_tiles;
                }
            }
            any_live = false;
#pragma unroll
            for (int s = 0; s < FLOW_SLOTS; ++s) any_live |= live[s];
        }
    } else if (warp > FLOW_MMA_WARP) {
        // ===================== feature gather (FLOW_PROD cp.async warps per slot) =====================
        // Row r of a tile = (point p0 + r / k, neighbour r % k) = [f_tmpl | f_src | dp hi,hi,lo | 0].  The two 128-byte
        // feature blocks of every row go global -> shared by 16-byte cp.async as soon as the slot's operand buffer is
        // free (the previous tile's last MMA has completed), i.e. while the epilogue warps still reduce that tile.
        // Random 128-byte rows of the L2-resident feature table arrive at ~15 B/cycle/SM whatever the lane mapping,
        // instruction or depth (tools/micro/lsu_gather.cu; TMA tile::gather4 manages 7, tools/micro/gather4.cu), so a
        // tile's 32 KB need ~2k cycles and the issuing warp is blocked for most of them: that is why the gather has
        // warp of its own.  One instruction covers 8 rows x 4 chunks: 8 lines for L1tex, and the 8 lanes of a
        // shared-memory store phase hit 8 different rows = all 32 banks.  Rows that do not exist are zero-filled.
        const int pw = warp - FLOW_MMA_WARP - 1, s = pw / FLOW_PROD, part = pw % FLOW_PROD;
        constexpr int RG = 128 / 8 / FLOW_PROD;                        // 8-row groups per gather warp
        const uint32_t act = ptx::smem_u32(s_act + s * act_bytes);
        const uint32_t stage = ptx::smem_u32(s_stage + pw * FLOW_STAGE_BYTES);    // one staging area per gather warp
        const int k = a.k, n = a.n, F = a.F, g = a.groups_per_tile, fc = F / 8, cq = lane >> 3;
        int prow[RG], pgrp[RG];
#pragma unroll
        for (int i = 0; i < RG; ++i) { prow[i] = part * (128 / FLOW_PROD) + i * 8 + (lane & 7); pgrp[i] = prow[i] / k; }
        // FC = chunks per feature block when known at compile time (0: runtime fc).  This warp is bound by its own
        // instruction issue, so the F = 64 case is fully unrolled (0.244 vs 0.261 ms per flow call with runtime loops).
        auto gather_tile = [&](auto fcc, const int b, const int p0, const int rows_valid) {
            constexpr int FC = decltype(fcc)::value;
            const int fcv = FC ? FC : fc;
            constexpr int CPL = FC ? (FC + 3) / 4 : 1;                  // chunks per lane and row (FC known)
            // template block: the g distinct template rows of the tile are fetched once into a staging area (every row
            // of a group repeats its template's 128 bytes) and replicated shared -> shared below
            const bool staged = FLOW_STAGE_BYTES > 0 && g * fcv * 16 <= FLOW_STAGE_BYTES;       // CTA-uniform
            if (staged) {
                for (int i = lane; i < g * fcv; i += 32) {
                    const int pr = i / fcv, c = i - pr * fcv;
                    const bool ok = p0 + pr < n;
                    ptx::cp_async16_cg(stage + i * 16,
                                       reinterpret_cast<const uint4 *>(a.feat + ((size_t)b * n + (ok ? p0 + pr : 0)) * F) + c,
                                       ok ? 16u : 0u);
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
            }
            // source block (and the template block when it is not staged): one gathered row per tile row
#pragma unroll
            for (int i = 0; i < RG; ++i) {
                const bool ok = prow[i] < rows_valid;
                const int src = ok ? __float_as_int(s_nbr[s * 128 + prow[i]].w) : 0;
                const uint4 *ftp = reinterpret_cast<const uint4 *>(a.feat + ((size_t)b * n + (ok ? p0 + pgrp[i] : 0)) * F);
                const uint4 *fsp = reinterpret_cast<const uint4 *>(a.feat + ((size_t)(a.B + b) * n + src) * F);
                const uint32_t d = act + prow[i] * 16, nb = ok ? 16u : 0u;
                if constexpr (FC > 0) {
#pragma unroll
                    for (int j = 0; j < CPL; ++j) {
                        const int c = cq + 4 * j;
                        if (FC % 4 == 0 || c < FC) {
                            if (!staged) ptx::cp_async16_ca(d + c * TC_LBO, ftp + c, nb);
                            ptx::cp_async16_cg(d + (FC + c) * TC_LBO, fsp + c, nb);
                        }
                    }
                } else {
                    for (int c = cq; c < fcv; c += 4) {
                        if (!staged) ptx::cp_async16_ca(d + c * TC_LBO, ftp + c, nb);
                        ptx::cp_async16_cg(d + (fcv + c) * TC_LBO, fsp + c, nb);
                    }
                }
            }
            if (staged) {
                asm volatile("cp.async.commit_group;" ::: "memory");
                asm volatile("cp.async.wait_group 1;" ::: "memory");    // my part of the staging area has landed
                __syncwarp();                                           // ... and every other lane's
                if constexpr (FC > 0) {
                    constexpr int BR = RG < 4 ? RG : 4;                 // row groups per batch: loads first, then stores
#pragma unroll
                    for (int i0 = 0; i0 < RG; i0 += BR) {
                        uint32_t v[BR][CPL][4];
#pragma unroll
                        for (int i = 0; i < BR; ++i)
#pragma unroll
                            for (int j = 0; j < CPL; ++j) {
                                const int c = cq + 4 * j;
                                v[i][j][0] = v[i][j][1] = v[i][j][2] = v[i][j][3] = 0u;
                                if (prow[i0 + i] < rows_valid && (FC % 4 == 0 || c < FC))
                                    ld_shared_v4(stage + (pgrp[i0 + i] * FC + c) * 16, v[i][j][0], v[i][j][1], v[i][j][2],
                                                 v[i][j][3]);
                            }
#pragma unroll
                        for (int i = 0; i < BR; ++i)
#pragma unroll
                            for (int j = 0; j < CPL; ++j) {
                                const int c = cq + 4 * j;
                                if (FC % 4 == 0 || c < FC)
                                    st_shared_v4(act + prow[i0 + i] * 16 + c * TC_LBO, v[i][j][0], v[i][j][1], v[i][j][2],
                                                 v[i][j][3]);
                            }
                    }
                } else {
#pragma unroll 1
                    for (int i = 0; i < RG; ++i) {
                        const bool ok = prow[i] < rows_valid;
                        for (int c = cq; c < fcv; c += 4) {
                            uint32_t v0 = 0u, v1 = 0u, v2 = 0u, v3 = 0u;
                            if (ok) ld_shared_v4(stage + (pgrp[i] * fcv + c) * 16, v0, v1, v2, v3);
                            st_shared_v4(act + prow[i] * 16 + c * TC_LBO, v0, v1, v2, v3);
                        }
                    }
                }
                __syncwarp();                                           // the staging area is reused by the next tile
            }
        };
        uint32_t iph = 0, fph = 0;
        const int tile0 = blockIdx.x * FLOW_SLOTS + s;
        for (int tile = tile0; tile < a.num_tiles; tile += tiles_per_iter) {
            const int b = tile / a.tiles_per_pair;
            const int p0 = (tile - b * a.tiles_per_pair) * g;
            const int rows_valid = min(g, n - p0) * k;
            ptx::mbar_wait(&idx_ready[s], iph & 1); ++iph;
            if (tile != tile0) { ptx::mbar_wait(&act_free[s], fph & 1); ++fph; }
            if (fc == 8) gather_tile(std::integral_constant<int, 8>{}, b, p0, rows_valid);   // F = 64: fully unrolled
            else gather_tile(std::integral_constant<int, 0>{}, b, p0, rows_valid);
            ptx::cp_async_wait_all();
            ptx::fence_proxy_async_smem();
            ptx::mbar_arrive(&gath[s]);
        }
    } else {
        // ===================== coordinates + epilogues (8 warps per slot) =====================
        // warp = (slot, column half, lane quarter): both halves of a quarter own the same 32 tile rows / TMEM lanes and
        // split the columns (hidden layers) and the 128-channel blocks (last layer).
        const int s = warp / FLOW_WPS, half = (warp % FLOW_WPS) >> 2, q = warp & 3;
        const int r = q * 32 + lane;                                   // tile row == TMEM lane
        const uint32_t act = ptx::smem_u32(s_act + s * act_bytes);
        const uint32_t act_row = act + r * 16;
        const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16) + s * 256;
        const int k = a.k, n = a.n, F = a.F, g = a.groups_per_tile;
        const int fc = F / 8;                                          // chunks per feature block
        uint32_t dphase = 0, iphase = 0, gphase = 0;
        [[maybe_unused]] int ft = -1;
        for (int tile = blockIdx.x * FLOW_SLOTS + s; tile < a.num_tiles; tile += tiles_per_iter) {
            ++ft;
            FLTRACE(0);
            const int b = tile / a.tiles_per_pair;
            const int p0 = (tile - b * a.tiles_per_pair) * g;
            const int rows_valid = min(g, n - p0) * k;
            // ---- what these warps add to the rows: displacement chunks + radius mask (thread = row, half 0) from the
            // row record in shared memory, zero K padding (half 1); then publish
            uint32_t *maskw = s_maskw + ((ft & 1) * FLOW_SLOTS + s) * 4;    // by tile parity: written here, read in the
                                                                       // last epilogue while fast warps may be a tile on
            ptx::mbar_wait(&idx_ready[s], iphase & 1); ++iphase;
            FLTRACE(11);
            if (half == 0) {                                           // warp-uniform
                bool masked = true;                                    // rows that do not exist never contribute
                if (r < rows_valid) {
                    const float4 rec = s_nbr[s * 128 + r];
                    const float dx = rec.x, dy = rec.y, dz = rec.z;
                    store_coord_chunks<T16>(act_row + (2 * fc) * TC_LBO, act_row + (2 * fc + 1) * TC_LBO, dx, dy, dz);
                    // radius mask in fp32 on the raw displacement (deepclr.py:220-222)
                    masked = a.radius > 0.f && sqrtf(dx * dx + dy * dy + dz * dz) >= a.radius;
                } else {
                    st_shared_v4(act_row + (2 * fc) * TC_LBO, 0u, 0u, 0u, 0u);
                    st_shared_v4(act_row + (2 * fc + 1) * TC_LBO, 0u, 0u, 0u, 0u);
                }
                const uint32_t mw = __ballot_sync(0xffffffffu, masked);
                if (lane == 0) maskw[q] = mw;
            } else {
                for (int c = 2 * fc + 2; c < K0p / 8; ++c) st_shared_v4(act_row + c * TC_LBO, 0u, 0u, 0u, 0u);
            }
            FLTRACE(8);
            ptx::mbar_wait(&gath[s], gphase & 1); ++gphase;            // the feature blocks have landed
            FLTRACE(9);
            ptx::fence_proxy_async_smem();
            FLTRACE(1);
            ptx::mbar_arrive(&ready[s]);
            // ---- layer 1 epilogue -> H1 (in place of the rows)
            ptx::mbar_wait(&done[s], dphase & 1); ++dphase;
            ptx::tc_fence_after();
            FLTRACE(2);
            epilogue_hidden<T16>(tlane, half * 64, half * 64 + 64, act_row);
            if (half == 1) store_ones_chunks<T16>(act_row, 128);       // layer 2's bias columns (K 128..143)
            ptx::tc_fence_before();
            ptx::fence_proxy_async_smem();
            FLTRACE(3);
            ptx::mbar_arrive(&ready[s]);
            // ---- layer 2 epilogue -> H2
            ptx::mbar_wait(&done[s], dphase & 1); ++dphase;
            ptx::tc_fence_after();
            FLTRACE(4);
            epilogue_hidden<T16>(tlane + 128, half * 64, half * 64 + 64, act_row);
            ptx::tc_fence_before();
            ptx::fence_proxy_async_smem();
            FLTRACE(5);
            ptx::mbar_arrive(&ready[s]);
            // ---- layer 3 epilogue (transposed): lane = output channel, columns = tile rows; max over each group
            ptx::mbar_wait(&done[s], dphase & 1); ++dphase;
            ptx::tc_fence_after();
            FLTRACE(6);
            uint32_t mk[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) mk[i] = maskw[i];
            for (int mb = half; mb < H3 / 128; mb += 2) {
                const int ch = mb * 128 + r;
                const float bias = s_bias[256 + ch];
                const size_t o0 = ((size_t)b * n + p0) * H3 + ch;
                uint16_t *out = a.out + o0;
                float *out32 = a.out32 ? a.out32 + o0 : nullptr;
                const uint32_t ta = tlane + mb * 128;
                if (k == 20) flow_group_max<20, T16>(ta, mk, k, g, n - p0, bias, out, out32, H3);
                else if (k == 30) flow_group_max<30, T16>(ta, mk, k, g, n - p0, bias, out, out32, H3);
                else flow_group_max<0, T16>(ta, mk, k, g, n - p0, bias, out, out32, H3);
            }
            ptx::tc_fence_before();
            FLTRACE(7);
        }
    }
    __syncthreads();
    if (warp == FLOW_MMA_WARP) ptx::tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------- output head
constexpr int HEAD_LAYERS = 5;
// Weight ring shape (overridable for tools/ab.sh).  Every tile costs an exposed mbarrier wait + commit (~115 cycles,
// tools/micro/mma_ring.cu), so tiles are as wide as the shared-memory budget allows: measured head time per 64 pairs
// 64 x 5 stages 0.184 ms | 80 x 4 0.174 | 96 x 3 0.168 | 96 x 2 0.192 | 128 x 2 0.177 | 112 x 3 0.163 (ONE issuing warp).
// With two issuing warps (below) stage s belongs to warp s % 2 (tile `it` of the stream sits in stage it % STAGES and
// belongs to warp it % 2, so STAGES must be even): a consumer then never waits on a stage whose previous occupant was
// another warp's, and the 1-bit phase parity of its mbarrier cannot alias.  6 x 12 KB.
// Measured with two issuing warps: 80 x 4 0.1427 ms | 64 x 4 0.151 | **48 x 6 0.140** | 32 x 10 0.160 | 112 x 2 0.185.
#ifndef DCLR_HEAD_KT
#define DCLR_HEAD_KT 48
#define DCLR_HEAD_STAGES 6
#endif
constexpr int HEAD_KT = DCLR_HEAD_KT;            // K elements per streamed weight tile (128 rows x 48 = 12 KB)
constexpr int HEAD_STAGES = DCLR_HEAD_STAGES;
constexpr int HEAD_TILE_BYTES = 128 * HEAD_KT * 2;
constexpr int HEAD_EPI_WARPS = 8;                // 4 TMEM lane quarters x 2 column halves
// TWO MMA-issuing warps: the issuing warp is bound by its own instruction latencies (~15 dependent scalar instructions per
// tcgen05.mma plus ~100 per weight tile: 85-100 cycles per 64-cycle MMA with one warp, tools/head_trace.py), not by the
// tensor pipe or the weight stream (copies disabled: same timeline).  Two warps with their own accumulator blocks keep
// the pipe fed (tools/micro/mma_dual.cu): warp q takes the output blocks nc = 2p + q of every layer, the weight tiles of
// a block pair are streamed interleaved (pair p: k-tile 0 of block 2p, k-tile 0 of block 2p+1, k-tile 1 of ...).
constexpr int HEAD_MMA_WARPS = 2;
// One producer lane per issuing warp: a lane needs ~345 cycles per cp.async.bulk whatever its size (wait + expect_tx +
// issue, tools/micro/tma_stream.cu), i.e. 59 B/cycle with 20 KB tiles against the 64 B/cycle the tensor pipe consumes.
constexpr int HEAD_PROD_WARP = HEAD_EPI_WARPS, HEAD_MMA_WARP = HEAD_EPI_WARPS + HEAD_MMA_WARPS;
constexpr int HEAD_THREADS = (HEAD_EPI_WARPS + 2 * HEAD_MMA_WARPS) * 32;
constexpr int HEAD_FIN_BUFS = 2 * HEAD_MMA_WARPS;    // last layer: TMEM blocks in flight, two per issuing warp
static_assert(HEAD_STAGES % HEAD_MMA_WARPS == 0 && HEAD_KT % 16 == 0, "weight ring: a stage has one consuming warp");

struct HeadArgs {
    const uint16_t *feat;   // [B*n, C0] motion features, 16-bit operand format
    const float *xyz;       // [B*n, 3] template sample coordinates
    const uint16_t *w;      // packed weight tiles in (layer, pair of 128-row blocks, K block, block of the pair) order
    const float *bias;      // biases of the 5 layers, concatenated
    float *pooled;          // [B, out[4]] zero-initialised; max over the points
    int *ticket;            // {next tile, CTAs finished}: both zero at launch, reset by the last CTA to leave
    int B, n, C0;
    int Kp[HEAD_LAYERS];    // padded input width per layer (multiple of 16)
    int out[HEAD_LAYERS];   // output width per layer (multiple of 128)
    int tiles_per_pair, num_tiles, act_k;        // act_k = widest activation (elements)
};

template <typename T16>
__global__ void __launch_bounds__(HEAD_THREADS, 1)
head_tc_kernel(HeadArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int nbias = 0;
    for (int l = 0; l < HEAD_LAYERS; ++l) nbias += a.out[l];
    uint8_t *s_act = smem;                                              // 128 x act_k x 2
    uint8_t *s_wt = s_act + (size_t)128 * a.act_k * 2;                  // HEAD_STAGES x HEAD_TILE_BYTES
    float *s_bias = reinterpret_cast<float *>(s_wt + HEAD_STAGES * HEAD_TILE_BYTES);
    uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_bias + nbias);
    uint64_t *full = s_bar, *empty = s_bar + HEAD_STAGES, *act_ready = empty + HEAD_STAGES,
             *layer_done = act_ready + 1, *fin_done = layer_done + 1, *fin_free = fin_done + HEAD_FIN_BUFS;
    uint64_t *next_full = fin_free + HEAD_FIN_BUFS;                      // [2]: s_next[i & 1] holds the tile of iteration i
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(next_full + 2);
    int *s_next = reinterpret_cast<int *>(s_tmem + 2);

    if (tid == 0) {
        ptx::mbar_init(&next_full[0], 1); ptx::mbar_init(&next_full[1], 1);
        for (int i = 0; i < HEAD_STAGES; ++i) { ptx::mbar_init(&full[i], 1); ptx::mbar_init(&empty[i], 1); }
        ptx::mbar_init(act_ready, HEAD_EPI_WARPS * 32);
        ptx::mbar_init(layer_done, HEAD_MMA_WARPS);
        for (int i = 0; i < HEAD_FIN_BUFS; ++i) { ptx::mbar_init(&fin_done[i], 1); ptx::mbar_init(&fin_free[i], HEAD_EPI_WARPS * 32); }
        ptx::fence_mbar_init();
    }
    if (warp == HEAD_MMA_WARP) ptx::tmem_alloc(s_tmem, 512);
    for (int e = tid; e < nbias; e += HEAD_THREADS) s_bias[e] = __ldg(a.bias + e);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *s_tmem;
    constexpr uint32_t IDESC = make_idesc(128, 128, Cvt<T16>::is_bf16);
    // Tiles are handed out by ticket: iteration 0 of a CTA is tile blockIdx.x, the tile of iteration i + 1 is drawn by
    // thread 0 while iteration i starts (atomicAdd on a.ticket) and published through s_next / next_full.  A CTA that
    // gets its SM late (the kernel shares the GPU with the next batch's FPS blocks, 128 SMs wide, on other streams)
    // then simply draws fewer tiles, instead of holding up the kernel with a static share.
    auto next_tile = [&](const int i) {                 // tile of iteration i + 1 (every role's lanes may call it)
        ptx::mbar_wait(&next_full[(i + 1) & 1], (i >> 1) & 1);
        return *reinterpret_cast<volatile int *>(&s_next[(i + 1) & 1]);
    };

    if (warp >= HEAD_PROD_WARP && warp < HEAD_MMA_WARP) {
        // ===================== weight producers: stream every tile of every layer, once per row tile ===========
        // producer `pq` feeds issuing warp `pq` (the tiles with it % HEAD_MMA_WARPS == pq, in that warp's stages)
        const int pq = warp - HEAD_PROD_WARP;
        if (lane == 0) {
            uint32_t it = 0;
            for (int ti = 0, tile = blockIdx.x; tile < a.num_tiles; tile = next_tile(ti), ++ti) {
                const uint8_t *src = reinterpret_cast<const uint8_t *>(a.w);
                for (int l = 0; l < HEAD_LAYERS; ++l) {
                    const int nblocks = a.out[l] / 128;
                    for (int p0 = 0; p0 < nblocks; p0 += HEAD_MMA_WARPS)
                        for (int k0 = 0; k0 < a.Kp[l]; k0 += HEAD_KT)
                            for (int q = 0; q < HEAD_MMA_WARPS; ++q, ++it) {
                            const uint32_t bytes = 128u * min(HEAD_KT, a.Kp[l] - k0) * 2;
                            if (q != pq) {
                                if (p0 + q < nblocks) src += bytes;
                                continue;
                            }
                            const int st = it % HEAD_STAGES;
                            ptx::mbar_wait(&empty[st], ((it / HEAD_STAGES) & 1) ^ 1);
                            if (p0 + q >= nblocks) {       // odd number of blocks: an empty slot keeps the stage <-> warp map
                                ptx::mbar_arrive(&full[st]);
                                continue;
                            }
#ifdef DCLR_HEAD_NOCOPY       // timing experiment only (tools/head_trace.py): the weights "arrive" at once
                            ptx::mbar_arrive(&full[st]);
#else
                            ptx::mbar_expect_tx(&full[st], bytes);
                            ptx::bulk_g2s(s_wt + st * HEAD_TILE_BYTES, src, bytes, &full[st]);
#endif
                            src += bytes;
                        }
                }
            }
        }
    } else if (warp >= HEAD_MMA_WARP) {
        // ===================== MMA issuers: each warp runs the loop converged, one elected lane issues ==========
        // Warp q owns the output blocks nc = 2p + q (and, in the last layer, the TMEM buffer q); `it` counts the weight
        // tiles of the stream, both warps count all of them and take their own.
        {
            const int q = warp - HEAD_MMA_WARP;
            const uint32_t act_lo = sdesc_lo(ptx::smem_u32(s_act)), wt_lo = sdesc_lo(ptx::smem_u32(s_wt));
            uint32_t it = 0, n_ready = 0, fin_use[2] = {0, 0};
            for (int ti = 0, tile = blockIdx.x; tile < a.num_tiles; tile = next_tile(ti), ++ti) {
                for (int l = 0; l < HEAD_LAYERS; ++l) {
                    ptx::mbar_wait(act_ready, n_ready & 1); ++n_ready;      // this layer's input is in s_act
                    ptx::tc_fence_after();
                    if (lane == 0 && q == 0) TRACE(ti * 64 + l * 2);
                    const bool last = l == HEAD_LAYERS - 1;
                    const int Kp = a.Kp[l], nblocks = a.out[l] / 128;
                    const int ktiles = (Kp + HEAD_KT - 1) / HEAD_KT;
                    for (int p0 = 0; p0 < nblocks; p0 += HEAD_MMA_WARPS) {
                        const int nc = p0 + q;
                        if (nc >= nblocks) {                           // empty slots of an odd block count: pass them on
                            uint32_t mine = it + q;
                            for (int k0 = 0; k0 < Kp; k0 += HEAD_KT, mine += HEAD_MMA_WARPS) {
                                const uint32_t st = mine % HEAD_STAGES;
                                ptx::mbar_wait(&full[st], (mine / HEAD_STAGES) & 1);
                                ptx::mma_commit_warp(&empty[st]);
                            }
                        } else {
                            uint32_t dcol = tmem_base + nc * 128;
                            const int buf = nc % HEAD_FIN_BUFS, bj = (nc / HEAD_MMA_WARPS) & 1;   // my two buffers alternate
                            if (last) {
                                if (fin_use[bj] > 0) {
                                    ptx::mbar_wait(&fin_free[buf], (fin_use[bj] - 1) & 1);
                                    ptx::tc_fence_after();
                                }
                                ++fin_use[bj];
                                dcol = tmem_base + buf * 128;
                            }
                            uint32_t a_lo = act_lo, acc = 0, mine = it + q;
                            for (int k0 = 0; k0 < Kp; k0 += HEAD_KT, mine += HEAD_MMA_WARPS) {
                                const uint32_t st = mine % HEAD_STAGES, ph = (mine / HEAD_STAGES) & 1;
                                ptx::mbar_wait(&full[st], ph);
                                ptx::tc_fence_after();
                                const uint32_t w_lo = wt_lo + st * (HEAD_TILE_BYTES >> 4);
                                // hidden layers: D = X.W^T (A = activations); last layer transposed (A = weights)
                                const uint32_t p_lo = last ? w_lo : a_lo, q_lo = last ? a_lo : w_lo;
                                if (Kp - k0 >= HEAD_KT) {                   // full K tile: HEAD_KT / 16 MMAs, unrolled
                                    ptx::mma_f16_ss_warp(dcol, sdesc(p_lo), sdesc(q_lo), IDESC, acc);
#pragma unroll
                                    for (int ks = 1; ks < HEAD_KT / 16; ++ks)
                                        ptx::mma_f16_ss_warp(dcol, sdesc(p_lo + ks * SDESC_KSTEP), sdesc(q_lo + ks * SDESC_KSTEP), IDESC, 1u);
                                } else {
                                    for (int ks = 0; ks < (Kp - k0) / 16; ++ks)
                                        ptx::mma_f16_ss_warp(dcol, sdesc(p_lo + ks * SDESC_KSTEP), sdesc(q_lo + ks * SDESC_KSTEP), IDESC, ks ? 1u : acc);
                                }
                                acc = 1;
                                a_lo += (HEAD_KT / 16) * SDESC_KSTEP;
                                ptx::mma_commit_warp(&empty[st]);      // frees the weight stage when the MMAs retire
                            }
                            if (last) ptx::mma_commit_warp(&fin_done[buf]);
                        }
                        it += ktiles * HEAD_MMA_WARPS;
                    }
                    if (!last) ptx::mma_commit_warp(layer_done);       // one arrival per issuing warp
                    if (lane == 0 && q == 0) TRACE(ti * 64 + l * 2 + 1);
                }
            }
        }
    } else {
        // ===================== row build + epilogues =====================
        // warp = (column half, lane quarter): both halves own the same 32 rows / TMEM lanes and split the columns
        const int half = warp >> 2, q = warp & 3;
        const int r = q * 32 + lane;
        const uint32_t act_row = ptx::smem_u32(s_act) + r * 16;
        const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16);
        uint32_t n_done = 0, fin_par = 0;            // bit b: phase parity of fin_done[b]
        const int C0 = a.C0, n = a.n;
        const int cc = C0 / 8, cmid = cc / 2, kc0 = a.Kp[0] / 8;
        const int cb = half == 0 ? 0 : cmid, ce = half == 0 ? cmid : cc;     // my feature chunks of a row
        // The NEXT tile's row half travels global -> registers while the last layer runs (these warps mostly wait then):
        // loaded at build time the 64 KB of a tile arrive through the LSU at ~15 B/cycle and cost 5-7 k cycles of every
        // tile (tools/head_trace.py).  HEAD_PRE chunks per thread, i.e. C0 = 256; other widths load at build time.
        constexpr int HEAD_PRE = 16;
        const bool can_pre = ce - cb == HEAD_PRE;
        uint4 pre[HEAD_PRE];
        float pre_xyz[3] = {0.f, 0.f, 0.f};
        auto prefetch_rows = [&](const int t) {
            const int tb = t / a.tiles_per_pair;
            const int tpt = (t - tb * a.tiles_per_pair) * 128 + r;
            if (tpt < n) {
                const size_t row = (size_t)tb * n + tpt;
                const uint4 *f = reinterpret_cast<const uint4 *>(a.feat + row * C0) + cb;
#pragma unroll
                for (int i = 0; i < HEAD_PRE; ++i) pre[i] = __ldg(f + i);
                if (half == 1) {
#pragma unroll
                    for (int i = 0; i < 3; ++i) pre_xyz[i] = __ldg(a.xyz + row * 3 + i);
                }
            }
        };
        if (can_pre && (int)blockIdx.x < a.num_tiles) prefetch_rows(blockIdx.x);
        int tile_next = 0;
        for (int ti = 0, tile = blockIdx.x; tile < a.num_tiles; tile = tile_next, ++ti) {
            const int b = tile / a.tiles_per_pair;
            const int p0 = (tile - b * a.tiles_per_pair) * 128;
            const int pt = p0 + r;
            [[maybe_unused]] const int tr0 = ti * 64 + 16;
            if (tid == 0) {
                TRACE(tr0);
                // draw the tile of iteration ti + 1; everybody has long read slot (ti + 1) & 1 (the tile of iteration
                // ti - 1, which is complete)
                s_next[(ti + 1) & 1] = (int)gridDim.x + atomicAdd(a.ticket, 1);
                ptx::mbar_arrive(&next_full[(ti + 1) & 1]);
            }
            // ---- input rows: [motion feat (C0) | xyz hi,hi,lo | 0]; half 0 writes the first feature chunks, half 1
            //      the rest, the coordinates and the padding
            if (pt < n) {
                const size_t row = (size_t)b * n + pt;
                if (can_pre) {
#pragma unroll
                    for (int i = 0; i < HEAD_PRE; ++i)
                        st_shared_v4(act_row + (cb + i) * TC_LBO, pre[i].x, pre[i].y, pre[i].z, pre[i].w);
                } else {
                    const uint4 *f = reinterpret_cast<const uint4 *>(a.feat + row * C0);
#pragma unroll 8
                    for (int c = cb; c < ce; ++c) {
                        const uint4 u = __ldg(f + c);
                        st_shared_v4(act_row + c * TC_LBO, u.x, u.y, u.z, u.w);
                    }
                }
                if (half == 1) {
                    if (!can_pre) {
#pragma unroll
                        for (int i = 0; i < 3; ++i) pre_xyz[i] = __ldg(a.xyz + row * 3 + i);
                    }
                    store_coord_chunks<T16>(act_row + cc * TC_LBO, act_row + (cc + 1) * TC_LBO, pre_xyz[0], pre_xyz[1],
                                            pre_xyz[2]);
                    for (int c = cc + 2; c < kc0; ++c) st_shared_v4(act_row + c * TC_LBO, 0u, 0u, 0u, 0u);
                }
            } else {
                for (int c = cb; c < ce; ++c) st_shared_v4(act_row + c * TC_LBO, 0u, 0u, 0u, 0u);
                if (half == 1)
                    for (int c = cc; c < kc0; ++c) st_shared_v4(act_row + c * TC_LBO, 0u, 0u, 0u, 0u);
            }
            ptx::fence_proxy_async_smem();
            if (tid == 0) TRACE(tr0 + 1);
            ptx::mbar_arrive(act_ready);
            // ---- hidden layers: accumulators -> activations, in place
            int boff = 0;
            for (int l = 0; l < HEAD_LAYERS - 1; ++l) {
                ptx::mbar_wait(layer_done, n_done & 1); ++n_done;
                ptx::tc_fence_after();
                if (tid == 0) TRACE(tr0 + 2 + l * 2);
                const int hw = a.out[l] / 2;
                epilogue_hidden<T16>(tlane, half * hw, half * hw + hw, act_row);
                if (half == 1 && l < HEAD_LAYERS - 2) store_ones_chunks<T16>(act_row, a.out[l]);   // next layer's bias
                boff += a.out[l];
                ptx::tc_fence_before();
                ptx::fence_proxy_async_smem();
                if (tid == 0) TRACE(tr0 + 3 + l * 2);
                ptx::mbar_arrive(act_ready);
            }
            tile_next = next_tile(ti);
            if (can_pre && tile_next < a.num_tiles) prefetch_rows(tile_next);
            // ---- last layer (transposed): lane = channel, columns = points of the tile; max over the valid points.
            //      The halves split the 128 columns; both contribute through atomicMax.
            const int nvalid = min(128, n - p0);
            for (int nc = 0; nc < a.out[HEAD_LAYERS - 1] / 128; ++nc) {
                const int buf = nc % HEAD_FIN_BUFS;
                ptx::mbar_wait(&fin_done[buf], (fin_par >> buf) & 1); fin_par ^= 1u << buf;
                ptx::tc_fence_after();
                if (tid == 0) TRACE(tr0 + 10 + nc);
                float cur = -INFINITY;
#pragma unroll
                for (int c0 = 0; c0 < 64; c0 += 32) {
                    uint32_t v[32];
                    ptx::tmem_ld32(tlane + buf * 128 + half * 64 + c0, v);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (half * 64 + c0 + i < nvalid) cur = fmaxf(cur, __uint_as_float(v[i]));
                }
                ptx::tc_fence_before();
                ptx::mbar_arrive(&fin_free[buf]);
                const int ch = nc * 128 + r;
                const float val = fmaxf(cur + s_bias[boff + ch], 0.f);
                // post-ReLU values are >= 0: their int image is order preserving
                if (half * 64 < nvalid)
                    atomicMax(reinterpret_cast<int *>(a.pooled) + (size_t)b * a.out[HEAD_LAYERS - 1] + ch,
                              __float_as_int(val));
            }
        }
    }
    __syncthreads();
    if (warp == HEAD_MMA_WARP) ptx::tmem_dealloc(tmem_base, 512);
    if (tid == 0 && atomicAdd(a.ticket + 1, 1) == (int)gridDim.x - 1) {     // every CTA has drawn its last ticket
        a.ticket[0] = 0;
        a.ticket[1] = 0;
        __threadfence();
    }
}

// ---------------------------------------------------------------- host side
struct TcModel {
    int fmt;                      // 1 bf16, 2 fp16
    int F, k, K0p, H3;
    float radius;
    uint16_t *flow_w = nullptr;   // packed flow weights
    float *flow_b = nullptr;
    uint16_t *head_w = nullptr;   // packed head weight tiles
    float *head_b = nullptr;
    int *tickets = nullptr;       // {next tile, CTAs finished} of head_tc_kernel: zero between launches
    int head_Kp[HEAD_LAYERS], head_out[HEAD_LAYERS], head_act_k;
    size_t flow_smem, head_smem;
    int num_sms;
};

static int round_up(int v, int m) { return (v + m - 1) / m * m; }

template <typename T16>
static int pack_rows(const float *W, const float *bias, int kbias, int out, int in, int nfeat, int row0, int R, int k0,
                     int Kt, uint16_t *dst) {
    pack_tile_kernel<T16><<<div_up(R * Kt, 256), 256>>>(W, bias, kbias, out, in, nfeat, row0, R, k0, Kt, dst);
    DCLR_LAUNCH_CHECK("pack_tile_kernel");
    return DCLR_OK;
}

static int pack_any(int fmt, const float *W, const float *bias, int kbias, int out, int in, int nfeat, int row0, int R,
                    int k0, int Kt,
                    uint16_t *dst) {
    return fmt == 1 ? pack_rows<__nv_bfloat16>(W, bias, kbias, out, in, nfeat, row0, R, k0, Kt, dst)
                    : pack_rows<__half>(W, bias, kbias, out, in, nfeat, row0, R, k0, Kt, dst);
}

#ifdef DCLR_TRACE
extern "C" int dclr_debug_read_flow_trace(long long *dst, int n) {
    return cudaMemcpyFromSymbol(dst, g_flow_trace, sizeof(long long) * (n < 1024 ? n : 1024)) == cudaSuccess ? 0 : -1;
}
extern "C" int dclr_debug_read_trace(long long *dst, int n) {
    return cudaMemcpyFromSymbol(dst, g_trace, sizeof(long long) * (n < 1024 ? n : 1024)) == cudaSuccess ? 0 : -1;
}
#endif

void tc_destroy(TcModel *t) {
    if (!t) return;
    if (t->flow_w) cudaFree(t->flow_w);
    if (t->flow_b) cudaFree(t->flow_b);
    if (t->head_w) cudaFree(t->head_w);
    if (t->head_b) cudaFree(t->head_b);
    if (t->tickets) cudaFree(t->tickets);
    delete t;
}

int tc_create(const TcLayerSpec *sp, int F, int k, float radius, int fmt, TcModel **out) {
    // shapes this path is built for (the reference model configs); anything else must use numerics fp32
    if (F % 8 || F < 8) return fail(DCLR_E_UNSUPPORTED, "tensor-core path: feature width %d is not a multiple of 8", F);
    if (sp[0].in != 2 * F + 3 || sp[0].out != 128 || sp[1].in != 128 || sp[1].out != 128 || sp[2].in != 128 ||
        sp[2].out % 128 || sp[2].out > 256)
        return fail(DCLR_E_UNSUPPORTED, "tensor-core path: merge MLP must be [%d,128,128,128|256]", 2 * F + 3);
    if (k < 1 || k > 128) return fail(DCLR_E_UNSUPPORTED, "tensor-core path: k must be in [1,128]");
    const int H3 = sp[2].out;
    if (sp[3].in != H3 + 3) return fail(DCLR_E_UNSUPPORTED, "tensor-core path: head input width mismatch");
    for (int l = 0; l < HEAD_LAYERS; ++l) {
        const TcLayerSpec &s = sp[3 + l];
        if (s.out % 128 || (l < HEAD_LAYERS - 1 && s.out > 512) || (l > 0 && (s.in != sp[2 + l].out || s.in % 16)))
            return fail(DCLR_E_UNSUPPORTED, "tensor-core path: head layer %d shape [%d -> %d] not supported", l, s.in,
                        s.out);
    }
    TcModel *t = new (std::nothrow) TcModel();
    if (!t) return fail(DCLR_E_INVALID, "out of host memory");
    t->fmt = fmt; t->F = F; t->k = k; t->radius = radius; t->H3 = H3;
    t->K0p = round_up(2 * F + 9 + 2, 16);          // features | 9 coordinate columns | 2 bias columns
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&t->num_sms, cudaDevAttrMultiProcessorCount, dev);
    int rc = DCLR_OK;
    auto bail = [&](int code) { tc_destroy(t); return code; };
#define TC_CUDA(expr)                                                                        \
    do {                                                                                     \
        cudaError_t e__ = (expr);                                                            \
        if (e__ != cudaSuccess)                                                              \
            return bail(fail(DCLR_E_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e__))); \
    } while (0)

    // ---- flow: W1 [128 x K0p] | W2 [128 x 128] | W3 in 128-row blocks [H3 x 128]; biases b1 b2 b3
    const size_t flow_elems = (size_t)128 * t->K0p + 128 * FLOW_K2P + (size_t)H3 * 128;
    TC_CUDA(cudaMalloc(&t->flow_w, flow_elems * 2));
    TC_CUDA(cudaMalloc(&t->flow_b, (256 + H3) * sizeof(float)));
    uint16_t *p = t->flow_w;
    if ((rc = pack_any(fmt, sp[0].W, sp[0].b, 2 * F + 9, 128, sp[0].in, 2 * F, 0, 128, 0, t->K0p, p))) return bail(rc);
    p += (size_t)128 * t->K0p;
    if ((rc = pack_any(fmt, sp[1].W, sp[1].b, 128, 128, 128, 0, 0, 128, 0, FLOW_K2P, p))) return bail(rc);
    p += 128 * FLOW_K2P;
    for (int mb = 0; mb < H3 / 128; ++mb, p += 128 * 128)
        if ((rc = pack_any(fmt, sp[2].W, nullptr, 0, H3, 128, 0, mb * 128, 128, 0, 128, p))) return bail(rc);
    TC_CUDA(cudaMemcpy(t->flow_b, sp[0].b, 128 * 4, cudaMemcpyDeviceToDevice));
    TC_CUDA(cudaMemcpy(t->flow_b + 128, sp[1].b, 128 * 4, cudaMemcpyDeviceToDevice));
    TC_CUDA(cudaMemcpy(t->flow_b + 256, sp[2].b, H3 * 4, cudaMemcpyDeviceToDevice));
    t->flow_smem = (size_t)128 * t->K0p * 2 + 128 * FLOW_K2P * 2 + (size_t)H3 * 128 * 2 +
                   (size_t)FLOW_SLOTS * 128 * t->K0p * 2 + (256 + H3) * 4 + 2 * FLOW_SLOTS * 4 * 4 +
                   (1 + 5 * FLOW_SLOTS) * 8 + 16 + FLOW_SLOTS * 128 * 16 + FLOW_SLOTS * FLOW_PROD * FLOW_STAGE_BYTES;

    // ---- head: tiles in (layer, pair of 128-row blocks, K block, block of the pair) order: head_tc_kernel's stream
    size_t head_elems = 0;
    int nbias = 0;
    t->head_act_k = 0;
    for (int l = 0; l < HEAD_LAYERS; ++l) {
        // hidden layers carry their bias as two extra K columns (l = 0: in the coordinate padding); the last layer adds
        // it in the epilogue, in fp32
        t->head_Kp[l] = l == 0 ? round_up(H3 + 9 + 2, 16) : (l < HEAD_LAYERS - 1 ? sp[3 + l].in + 16 : sp[3 + l].in);
        t->head_out[l] = sp[3 + l].out;
        head_elems += (size_t)t->head_out[l] * t->head_Kp[l];
        nbias += t->head_out[l];
        if (t->head_Kp[l] > t->head_act_k) t->head_act_k = t->head_Kp[l];
        if (l < HEAD_LAYERS - 1 && t->head_out[l] > t->head_act_k) t->head_act_k = t->head_out[l];
    }
    TC_CUDA(cudaMalloc(&t->head_w, head_elems * 2));
    TC_CUDA(cudaMalloc(&t->head_b, nbias * sizeof(float)));
    TC_CUDA(cudaMalloc(&t->tickets, 4 * sizeof(int)));
    TC_CUDA(cudaMemset(t->tickets, 0, 4 * sizeof(int)));
    p = t->head_w;
    int boff = 0;
    for (int l = 0; l < HEAD_LAYERS; ++l) {
        const TcLayerSpec &s = sp[3 + l];
        for (int p0 = 0; p0 < s.out / 128; p0 += HEAD_MMA_WARPS)
            for (int k0 = 0; k0 < t->head_Kp[l]; k0 += HEAD_KT)
                for (int nc = p0; nc < std::min(p0 + HEAD_MMA_WARPS, s.out / 128); ++nc) {
                    const int kt = t->head_Kp[l] - k0 < HEAD_KT ? t->head_Kp[l] - k0 : HEAD_KT;
                    if ((rc = pack_any(fmt, s.W, l < HEAD_LAYERS - 1 ? s.b : nullptr, l == 0 ? H3 + 9 : s.in, s.out, s.in,
                                       l == 0 ? H3 : 0, nc * 128, 128, k0, kt, p)))
                        return bail(rc);
                    p += (size_t)128 * kt;
                }
        TC_CUDA(cudaMemcpy(t->head_b + boff, s.b, s.out * 4, cudaMemcpyDeviceToDevice));
        boff += s.out;
    }
    t->head_smem = (size_t)128 * t->head_act_k * 2 + HEAD_STAGES * HEAD_TILE_BYTES + (size_t)nbias * 4 +
                   (2 * HEAD_STAGES + 2 + 2 * HEAD_FIN_BUFS + 2) * 8 + 32;
    TC_CUDA(cudaDeviceSynchronize());
    if (t->flow_smem > 232448 || t->head_smem > 232448)
        return bail(fail(DCLR_E_UNSUPPORTED, "tensor-core path: shared memory budget exceeded (%zu / %zu bytes)",
                         t->flow_smem, t->head_smem));
    if (fmt == 1) {
        TC_CUDA(cudaFuncSetAttribute(flow_tc_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)t->flow_smem));
        TC_CUDA(cudaFuncSetAttribute(head_tc_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)t->head_smem));
    } else {
        TC_CUDA(cudaFuncSetAttribute(flow_tc_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)t->flow_smem));
        TC_CUDA(cudaFuncSetAttribute(head_tc_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)t->head_smem));
    }
#undef TC_CUDA
    *out = t;
    return DCLR_OK;
}

template <typename T16>
__global__ void convert16_kernel(const float *__restrict__ src, uint16_t *__restrict__ dst, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        dst[i] = Cvt<T16>::one(__ldg(src + i));
}

int tc_convert(TcModel *t, const float *src, uint16_t *dst, size_t n, cudaStream_t st) {
    const int grid = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
    if (t->fmt == 1) convert16_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(src, dst, n);
    else convert16_kernel<__half><<<grid, 256, 0, st>>>(src, dst, n);
    DCLR_LAUNCH_CHECK("convert16_kernel");
    return DCLR_OK;
}

int tc_flow(TcModel *t, const uint16_t *feat, const float4 *nbr, int B, int n, uint16_t *motion, float *motion32,
            cudaStream_t st) {
    FlowArgs fa;
    fa.feat = feat; fa.nbr = nbr; fa.w = t->flow_w; fa.bias = t->flow_b; fa.out = motion;
    fa.out32 = motion32;
    fa.B = B; fa.n = n; fa.k = t->k; fa.F = t->F; fa.K0p = t->K0p; fa.H3 = t->H3; fa.radius = t->radius;
    fa.groups_per_tile = 128 / t->k;
    fa.tiles_per_pair = div_up(n, fa.groups_per_tile);
    fa.num_tiles = B * fa.tiles_per_pair;
    const int grid = std::min(t->num_sms, div_up(fa.num_tiles, FLOW_SLOTS));
    if (t->fmt == 1) flow_tc_kernel<__nv_bfloat16><<<grid, FLOW_THREADS, t->flow_smem, st>>>(fa);
    else flow_tc_kernel<__half><<<grid, FLOW_THREADS, t->flow_smem, st>>>(fa);
    DCLR_LAUNCH_CHECK("flow_tc_kernel");
    return DCLR_OK;
}

int tc_head(TcModel *t, const float *xyz, const uint16_t *motion, int B, int n, float *pooled, cudaStream_t st) {
    HeadArgs ha;
    ha.feat = motion; ha.xyz = xyz; ha.w = t->head_w; ha.bias = t->head_b; ha.pooled = pooled; ha.ticket = t->tickets;
    ha.B = B; ha.n = n; ha.C0 = t->H3;
    for (int l = 0; l < HEAD_LAYERS; ++l) { ha.Kp[l] = t->head_Kp[l]; ha.out[l] = t->head_out[l]; }
    ha.tiles_per_pair = div_up(n, 128);
    ha.num_tiles = B * ha.tiles_per_pair;
    ha.act_k = t->head_act_k;
    const int grid = std::min(t->num_sms, ha.num_tiles);
    DCLR_CUDA(cudaMemsetAsync(pooled, 0, (size_t)B * t->head_out[HEAD_LAYERS - 1] * sizeof(float), st));
    if (t->fmt == 1) head_tc_kernel<__nv_bfloat16><<<grid, HEAD_THREADS, t->head_smem, st>>>(ha);
    else head_tc_kernel<__half><<<grid, HEAD_THREADS, t->head_smem, st>>>(ha);
    DCLR_LAUNCH_CHECK("head_tc_kernel");
    return DCLR_OK;
}

int tc_format(const TcModel *t) { return t->fmt; }

}  // namespace dclr
