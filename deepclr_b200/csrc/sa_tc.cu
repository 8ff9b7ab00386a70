// Set-abstraction grouping + shared MLP + max-pool on the tcgen05 tensor cores (dense neighbourhoods).  Same contract as
// sa_mlp.cu (pointnet2_utils.py:249-257, pytorch_utils.py:21-32, pointnet2_modules.py:40-52; padded duplicate rows are
// skipped exactly), same packing of the genuine (centre, neighbour) rows, but the three 1x1 convolutions of a 128-row tile
// are tensor-core GEMMs with the accumulators in TMEM.
//
// Precision: fp32-class, in every numerics mode.  An fp32 value x travels as TWO fp16 numbers, hi = fp16(x) and
// lo = fp16(x - hi) (22 significant bits), and a product w * x is evaluated as w_hi x_hi + w_hi x_lo + w_lo x_hi by
// laying the terms out along K:
//     activations  [ x_hi | x_lo | x_hi | 1 1 0 .. ]      weights  [ w_hi | w_hi | w_lo | b_hi b_lo 0 .. ]
// (the bias rides in the GEMM against two columns of ones).  Accumulation is fp32 in TMEM.  The inputs are tiny
// (C_in = 4 -> 16 -> 16 -> 32), so tripling K costs nothing: layer 1 is ONE 128 x 16 x 16 instruction, layers 2 and 3
// four each.  What the SIMT kernel spends per row -- 832 FFMA + ~750 other instructions, 23 % of the FFMA peak at every
// density (profiles/r2_sa_density_sweep.jsonl) -- becomes ~300 instructions of row building and epilogue.
//
// One CTA (4 warps) owns 32 consecutive centres of a cloud and walks their rows in tiles of 128, scale after scale:
//   build   thread = row: (centre, j) by binary search in the prefix sums, gather the neighbour, subtract the centre,
//           split, two 16-byte stores into the K-major operand tile (canonical no-swizzle UMMA layout, see mlp_tc.cu)
//   layer 1 D[rows, 16] = X W1'^T            -> epilogue: ReLU, split, six 16-byte stores (in place)
//   layer 2 D[rows, 16] = H1 W2'^T           -> same
//   layer 3 D[rows, 32] = H2 W3'^T           -> epilogue: every warp turns its 32 rows x 32 channels through shared memory
//           (the operand tile is free by then; 16-byte swizzle, conflict free both ways) so that lane = channel, keeps a
//           running max over the rows of a centre and flushes each run into a [32 centres][32] shared array with
//           atomicMax on the integer image (post-ReLU values are >= 0; order independent => deterministic)
// Everything inside a CTA is sequential (one __syncthreads + one mbarrier wait per layer); EIGHT CTAs per SM (64 TMEM
// columns and 28 KB of shared memory each) overlap each other's phases.  (A first version computed layer 3 transposed
// with the weights replicated over the four lane quarters: 128 TMEM columns and 40 KB per CTA, four CTAs per SM, 25 % of
// the warp slots -- 3.6 ms per 2048 ModelNet-shaped clouds against 4.2 ms SIMT.)
#include "common.cuh"
#include "kernels.cuh"
#include "ptx.cuh"
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <algorithm>

namespace dclr {

namespace {

constexpr int ST_CB = 32;                 // centres per CTA
constexpr int ST_ROWS = 128;              // rows per tile
constexpr int ST_K1 = 16, ST_K2 = 64;     // padded K of layer 1 / layers 2 and 3
constexpr int ST_LBO = ST_ROWS * 16;      // bytes between K chunks of a 128-row tile
constexpr int ST_W1 = 16 * ST_K1, ST_W2 = 16 * ST_K2, ST_W3 = 32 * ST_K2;       // halfs per packed image
constexpr int ST_WHALFS = ST_W1 + ST_W2 + ST_W3;                                // per scale: 3328 halfs = 6656 bytes
constexpr int ST_THREADS = 128;
#ifndef DCLR_SA_TC_CTAS
#define DCLR_SA_TC_CTAS 8
#endif
constexpr int ST_CTAS_PER_SM = DCLR_SA_TC_CTAS;

__host__ __device__ constexpr uint32_t st_idesc(int M, int N) {     // kind::f16, fp16 operands, fp32 accumulate, K-major
    return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ uint64_t st_sdesc(uint32_t smem_addr, uint32_t lbo_bytes) {
    return (uint64_t)((smem_addr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
           ((uint64_t)(128u >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ void st_sts4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void st_tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
// (a, b) -> packed fp16 pairs hi = {fp16(a), fp16(b)} (saturating) and lo = {fp16(a - hi_a), fp16(b - hi_b)}
__device__ __forceinline__ void st_split2(float a, float b, uint32_t &hi, uint32_t &lo) {
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(b), "f"(a));
    const float2 back = __half22float2(*reinterpret_cast<const __half2 *>(&hi));
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(b - back.y), "f"(a - back.x));
}
// the same after a ReLU of both values: max(., 0) is part of the first conversion
__device__ __forceinline__ void st_split2_relu(float a, float b, uint32_t &hi, uint32_t &lo) {
    asm("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(b), "f"(a));
    const float2 back = __half22float2(*reinterpret_cast<const __half2 *>(&hi));
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(fmaxf(b, 0.f) - back.y), "f"(fmaxf(a, 0.f) - back.x));
}

struct SaTcParams {
    const float *x;          // [B,N,CIN]
    const float *new_xyz;    // [B,M,3]
    const int32_t *idx[2];   // compact neighbour lists [B,M,ns]
    const int32_t *cnt[2];   // [B,M]
    int ns[2];
    const uint16_t *w;       // packed fp16 images, per scale: W1' [16 x 16] | W2' [16 x 64] | W3' [32 x 64]
    float *out;              // [B,M,ld_out]
    uint16_t *out16;         // or null
    int N, M, ld_out, nscales, fmt16, cin;
};

}  // namespace

__global__ void __launch_bounds__(ST_THREADS, ST_CTAS_PER_SM)
sa_tc_kernel(const SaTcParams a) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *s_act = smem;                                            // [128 rows x 64 K] fp16, 16 KB
    uint8_t *s_w = s_act + ST_ROWS * ST_K2 * 2;                       // this scale's weight images, 6.5 KB
    int *s_out = reinterpret_cast<int *>(s_w + ST_WHALFS * 2);        // [32 centres][32 channels], integer image of the max
    float *s_ctr = reinterpret_cast<float *>(s_out + ST_CB * 32);     // [32][3]
    uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_ctr + ST_CB * 3);  // [0] weights landed, [1] MMAs done
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(s_bar + 2);
    int *s_excl = reinterpret_cast<int *>(s_tmem + 2);                // [33] prefix sums of the rows per centre
    int *s_nraw = s_excl + ST_CB + 1;                                 // [32] hits per centre (0 = empty ball)
    uint8_t *s_ci = reinterpret_cast<uint8_t *>(s_nraw + ST_CB);      // [128] centre of every row of the current tile

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.y, c0 = blockIdx.x * ST_CB;
    const float *xb = a.x + (size_t)b * a.N * a.cin;

    if (tid == 0) {
        ptx::mbar_init(&s_bar[0], 1);
        ptx::mbar_init(&s_bar[1], 1);
        ptx::fence_mbar_init();
    }
    if (warp == 0) ptx::tmem_alloc(s_tmem, 64);
    if (tid < ST_CB * 3) {
        const int c = min(c0 + tid / 3, a.M - 1);
        s_ctr[tid] = __ldg(a.new_xyz + ((size_t)b * a.M + c) * 3 + tid % 3);
    }
    for (int e = tid; e < ST_CB * 32; e += ST_THREADS) s_out[e] = 0;
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *s_tmem;
    const uint32_t tlane = tmem_base + ((uint32_t)(warp * 32) << 16);
    const uint32_t act_a = ptx::smem_u32(s_act), act_row = act_a + tid * 16;
    const uint32_t w1_a = ptx::smem_u32(s_w), w2_a = w1_a + ST_W1 * 2, w3_a = w2_a + ST_W2 * 2;
    constexpr uint32_t ID16 = st_idesc(128, 16), ID32 = st_idesc(128, 32);
    uint32_t mma_phase = 0;
    // warp 0 issues (converged, one elected lane per instruction: ptx.cuh), everybody waits for the completion
    auto run_mma = [&](auto issue) {
        ptx::tc_fence_before();
        ptx::fence_proxy_async_smem();
        __syncthreads();
        if (warp == 0) {
            ptx::tc_fence_after();
            issue();
            ptx::mma_commit_warp(&s_bar[1]);
        }
        ptx::mbar_wait(&s_bar[1], mma_phase & 1);
        ++mma_phase;
        ptx::tc_fence_after();
    };
    // accumulators [16] -> ReLU -> (hi, lo) -> K chunks [hi 0-7][hi 8-15][lo 0-7][lo 8-15][hi 0-7][hi 8-15] of this row
    auto hidden_epilogue = [&](uint32_t taddr) {
        uint32_t v[16];
        st_tmem_ld16(taddr, v);
        ptx::tmem_ld_wait();
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) st_split2_relu(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]), hi[i], lo[i]);
        st_sts4(act_row + 0 * ST_LBO, hi[0], hi[1], hi[2], hi[3]);
        st_sts4(act_row + 1 * ST_LBO, hi[4], hi[5], hi[6], hi[7]);
        st_sts4(act_row + 2 * ST_LBO, lo[0], lo[1], lo[2], lo[3]);
        st_sts4(act_row + 3 * ST_LBO, lo[4], lo[5], lo[6], lo[7]);
        st_sts4(act_row + 4 * ST_LBO, hi[0], hi[1], hi[2], hi[3]);
        st_sts4(act_row + 5 * ST_LBO, hi[4], hi[5], hi[6], hi[7]);
    };

    for (int s = 0; s < a.nscales; ++s) {
        // ---- this scale's weights (one bulk copy) and the row bookkeeping of the 32 centres
        if (tid == 0) {
            ptx::mbar_expect_tx(&s_bar[0], ST_WHALFS * 2);
            ptx::bulk_g2s(s_w, a.w + (size_t)s * ST_WHALFS, ST_WHALFS * 2, &s_bar[0]);
        }
        if (warp == 1) {
            const int c = c0 + lane;
            const int n_raw = c < a.M ? __ldg(a.cnt[s] + (size_t)b * a.M + c) : 0;
            const int n_rows = c < a.M ? (n_raw > 0 ? n_raw : 1) : 0;          // empty ball: all slots are index 0
            int incl = n_rows;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int o = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += o;
            }
            s_excl[lane] = incl - n_rows;
            s_nraw[lane] = n_raw;
            if (lane == 31) s_excl[32] = incl;
        }
        __syncthreads();
        ptx::mbar_wait(&s_bar[0], s & 1);
        const int total = s_excl[ST_CB];
        const int32_t *lists = a.idx[s] + ((size_t)b * a.M + c0) * a.ns[s];
        // row `row` of this scale: its centre and the neighbour's point, fetched one tile ahead (index, then the point it
        // names: two dependent L2 round trips that would otherwise sit in front of every tile).
        // Two stages, because a warp issues in order: the index load's first USE (the point address) would stall it for
        // an L2 round trip.  fetch_index right after a tile's build, fetch_point one layer later, the point's first use
        // at the next build.
        int nci = 0, nk = 0;
        float nin[4] = {0.f, 0.f, 0.f, 0.f};
        auto fetch_index = [&](const int row) {
            nk = 0;
            if (row < total) {
                int ci = 0;
#pragma unroll
                for (int step = 16; step >= 1; step >>= 1)
                    if (s_excl[ci + step] <= row) ci += step;          // last centre whose first row is <= row
                const int j = row - s_excl[ci];
                if (s_nraw[ci] > 0) nk = __ldg(lists + (size_t)ci * a.ns[s] + j);
                nci = ci;
            } else {
                nci = 0xff;
            }
        };
        auto fetch_point = [&]() {
            if (nci == 0xff) return;
            if (a.cin == 4) {
                const float4 v = __ldg(reinterpret_cast<const float4 *>(xb) + nk);
                nin[0] = v.x; nin[1] = v.y; nin[2] = v.z; nin[3] = v.w;
            } else {
                nin[0] = __ldg(xb + (size_t)nk * 3); nin[1] = __ldg(xb + (size_t)nk * 3 + 1);
                nin[2] = __ldg(xb + (size_t)nk * 3 + 2); nin[3] = 0.f;
            }
        };
        fetch_index(tid);
        fetch_point();
        for (int r0 = 0; r0 < total; r0 += ST_ROWS) {
            // ---- build: this thread's row (fetched during the previous tile)
            const int ci = nci;
            {
                uint32_t w0 = 0u, w1 = 0u, w2 = 0u, w3 = 0u, ones = 0u;
                if (ci != 0xff) {
                    // grouped_xyz -= new_xyz (pointnet2_utils.py:252)
                    st_split2(nin[0] - s_ctr[ci * 3], nin[1] - s_ctr[ci * 3 + 1], w0, w2);
                    st_split2(nin[2] - s_ctr[ci * 3 + 2], nin[3], w1, w3);
                    ones = 0x3c003c00u;
                }
                st_sts4(act_row + 0 * ST_LBO, w0, w1, w2, w3);          // [x_hi (4) | x_lo (4)]
                st_sts4(act_row + 1 * ST_LBO, w0, w1, ones, 0u);        // [x_hi (4) | 1 1 0 0]
                // K chunks 6 (ones against the bias columns of layers 2 and 3) and 7 (zeros): the last epilogue of
                // the previous tile used the whole tile as its transpose buffer
                st_sts4(act_row + 6 * ST_LBO, 0x3c003c00u, 0u, 0u, 0u);
                st_sts4(act_row + 7 * ST_LBO, 0u, 0u, 0u, 0u);
            }
            // runs of equal centre among this warp's 32 rows: bit i = row i starts a run (row 0 always does)
            const int ci_up = __shfl_up_sync(0xffffffffu, ci, 1);
            const unsigned starts = __ballot_sync(0xffffffffu, lane == 0 || ci != ci_up);
            s_ci[tid] = (uint8_t)ci;
            fetch_index(r0 + ST_ROWS + tid);
            // ---- layer 1: D[rows, 16] at TMEM columns 0..15
            run_mma([&] { ptx::mma_f16_ss_warp(tmem_base, st_sdesc(act_a, ST_LBO), st_sdesc(w1_a, 16 * 16), ID16, 0u); });
            hidden_epilogue(tlane);
            fetch_point();                       // the index fetched after the build has arrived by now
            // ---- layer 2: columns 16..31
            run_mma([&] {
#pragma unroll
                for (int ks = 0; ks < ST_K2 / 16; ++ks)
                    ptx::mma_f16_ss_warp(tmem_base + 16, st_sdesc(act_a + ks * 2 * ST_LBO, ST_LBO),
                                         st_sdesc(w2_a + ks * 2 * 16 * 16, 16 * 16), ID16, ks > 0);
            });
            hidden_epilogue(tlane + 16);
            // ---- layer 3: D[rows, 32] at columns 32..63
            run_mma([&] {
#pragma unroll
                for (int ks = 0; ks < ST_K2 / 16; ++ks)
                    ptx::mma_f16_ss_warp(tmem_base + 32, st_sdesc(act_a + ks * 2 * ST_LBO, ST_LBO),
                                         st_sdesc(w3_a + ks * 2 * 32 * 16, 32 * 16), ID32, ks > 0);
            });
            {
                // this warp's 32 rows x 32 channels through its 4 KB of the (now free) operand tile: row r at byte
                // 128 r, its eight 16-byte groups swizzled by r & 7 -> thread = row writes, lane = channel reads
                uint32_t v[32];
                ptx::tmem_ld32(tlane + 32, v);
                ptx::tmem_ld_wait();
                const uint32_t trow = act_a + tid * 128;
#pragma unroll
                for (int g4 = 0; g4 < 8; ++g4)
                    st_sts4(trow + ((g4 ^ (tid & 7)) << 4), v[4 * g4], v[4 * g4 + 1], v[4 * g4 + 2], v[4 * g4 + 3]);
                __syncwarp();
                const float *tbuf = reinterpret_cast<const float *>(s_act) + warp * 32 * 32;
                float cur = 0.f;                                        // ReLU: the max starts at 0
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    cur = fmaxf(cur, tbuf[i * 32 + ((((lane >> 2) ^ (i & 7)) << 2) | (lane & 3))]);
                    if (i == 31 || ((starts >> (i + 1)) & 1u)) {        // warp-uniform: row i ends a run
                        const int c = s_ci[warp * 32 + i];              // broadcast read
                        if (c != 0xff) atomicMax(&s_out[c * 32 + lane], __float_as_int(cur));
                        cur = 0.f;
                    }
                }
            }
            ptx::tc_fence_before();
            __syncthreads();       // TMEM columns and the operand tile are free again; s_ci may be rewritten
        }
        // ---- the scale's 32 x 32 results: 4 centres per pass, 128 bytes contiguous per centre
        __syncthreads();
        for (int e = tid; e < ST_CB * 32; e += ST_THREADS) {
            const int ci = e >> 5, ch = e & 31;
            if (c0 + ci < a.M) {
                const float val = __int_as_float(s_out[e]);
                const size_t o = ((size_t)b * a.M + c0 + ci) * a.ld_out + s * 32 + ch;
                a.out[o] = val;
                if (a.out16)
                    a.out16[o] = a.fmt16 == 1 ? __bfloat16_as_ushort(__float2bfloat16_rn(val)) : __half_as_ushort(__float2half_rn(val));
            }
            s_out[e] = 0;
        }
        __syncthreads();
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc(tmem_base, 64);
}

// ---------------------------------------------------------------- host side
size_t sa_tc_weight_bytes() { return (size_t)2 * ST_WHALFS * sizeof(uint16_t); }

static void split_h(float x, uint16_t &hi, uint16_t &lo) {
    const float c = x > 65504.f ? 65504.f : (x < -65504.f ? -65504.f : x);
    const __half h = __float2half_rn(c);
    const __half l = __float2half_rn(x - __half2float(h));
    hi = __half_as_ushort(h);
    lo = __half_as_ushort(l);
}

// w_host: W1[16*cin] b1[16] W2[16*16] b2[16] W3[32*16] b3[32] (fp32, row-major [out,in]) -> the three packed images of
// one scale.  Image element (row n, column k) of an R-row tile sits at (k / 8) * (R * 8) + n * 8 + (k % 8).
void sa_tc_pack_scale(const float *w_host, int cin, uint16_t *img) {
    const float *W1 = w_host, *B1 = W1 + 16 * cin, *W2 = B1 + 16, *B2 = W2 + 16 * 16, *W3 = B2 + 16, *B3 = W3 + 32 * 16;
    for (int i = 0; i < ST_WHALFS; ++i) img[i] = 0;
    auto at = [](uint16_t *base, int R, int n, int k) -> uint16_t & { return base[(size_t)(k / 8) * (R * 8) + n * 8 + (k % 8)]; };
    uint16_t *i1 = img, *i2 = img + ST_W1, *i3 = i2 + ST_W2;
    for (int o = 0; o < 16; ++o) {                 // [w_hi (4) | w_hi (4) | w_lo (4) | b_hi b_lo 0 0]
        for (int i = 0; i < cin; ++i) {
            uint16_t hi, lo;
            split_h(W1[o * cin + i], hi, lo);
            at(i1, 16, o, i) = hi; at(i1, 16, o, 4 + i) = hi; at(i1, 16, o, 8 + i) = lo;
        }
        uint16_t bh, bl;
        split_h(B1[o], bh, bl);
        at(i1, 16, o, 12) = bh; at(i1, 16, o, 13) = bl;
    }
    auto pack64 = [&](uint16_t *im, int R, int row, const float *wrow, float bias) {   // [w_hi 16 | w_hi 16 | w_lo 16 | b_hi b_lo ..]
        for (int i = 0; i < 16; ++i) {
            uint16_t hi, lo;
            split_h(wrow[i], hi, lo);
            at(im, R, row, i) = hi; at(im, R, row, 16 + i) = hi; at(im, R, row, 32 + i) = lo;
        }
        uint16_t bh, bl;
        split_h(bias, bh, bl);
        at(im, R, row, 48) = bh; at(im, R, row, 49) = bl;
    };
    for (int o = 0; o < 16; ++o) pack64(i2, 16, o, W2 + o * 16, B2[o]);
    for (int o = 0; o < 32; ++o) pack64(i3, 32, o, W3 + o * 16, B3[o]);
}

int sa_tc_launch(const SaArgs &s, const uint16_t *w_dev, int B, int cin, cudaStream_t st) {
    SaTcParams p;
    p.x = s.x; p.new_xyz = s.new_xyz; p.out = s.out; p.out16 = s.out16; p.fmt16 = s.fmt16; p.N = s.N; p.M = s.M;
    p.ld_out = s.ld_out; p.nscales = s.nscales; p.cin = cin; p.w = w_dev;
    for (int i = 0; i < 2; ++i) { p.idx[i] = s.idx[i]; p.cnt[i] = s.cnt[i]; p.ns[i] = s.ns[i]; }
    // shared memory: the carve-up of the kernel; eight CTAs (8 x 64 TMEM columns) per SM
    const size_t smem = (size_t)ST_ROWS * ST_K2 * 2 + ST_WHALFS * 2 + ST_CB * 32 * 4 + ST_CB * 3 * 4 + 2 * 8 + 2 * 4 +
                        (ST_CB + 1) * 4 + ST_CB * 4 + ST_ROWS + 16;
    static_assert(ST_CTAS_PER_SM * ((size_t)ST_ROWS * ST_K2 * 2 + ST_WHALFS * 2 + ST_CB * 32 * 4 + 1024 + 1024) <= 233472,
                  "CTAs per SM: shared memory");
    static bool attr_done[64];
    int dev = 0;
    DCLR_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_done[dev]) {
        DCLR_CUDA(cudaFuncSetAttribute(sa_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_done[dev] = true;
    }
    sa_tc_kernel<<<dim3(div_up(s.M, ST_CB), B), ST_THREADS, smem, st>>>(p);
    DCLR_LAUNCH_CHECK("sa_tc_kernel");
    return DCLR_OK;
}

// rows[0] += genuine rows of this launch's centres (hits, or 1 for an empty ball): the density probe of model.cu
__global__ void sa_count_rows_kernel(const int32_t *__restrict__ cnt, size_t n, unsigned long long *rows) {
    unsigned long long acc = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int c = __ldg(cnt + i);
        acc += (unsigned long long)(c > 0 ? c : 1);
    }
    for (int d = 16; d >= 1; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if ((threadIdx.x & 31) == 0 && acc) atomicAdd(rows, acc);
}

int sa_count_rows_launch(const int32_t *cnt, size_t n, unsigned long long *rows, cudaStream_t st) {
    sa_count_rows_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 296), 256, 0, st>>>(cnt, n, rows);
    DCLR_LAUNCH_CHECK("sa_count_rows_kernel");
    return DCLR_OK;
}

}  // namespace dclr
