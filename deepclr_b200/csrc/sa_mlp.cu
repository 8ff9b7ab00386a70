// Set-abstraction grouping + shared MLP + max-pool, fused (fp32 SIMT).  Replaces, per scale, grouping_operation x2 +
// cat + 3x(Conv2d 1x1 + bias + ReLU) + max_pool2d of the reference (pointnet2_utils.py:249-257,
// pytorch_utils.py:21-32, pointnet2_modules.py:40-52): the grouped tensor [B, 3+C, npoint, nsample] is never formed.
//
// Exactness of skipping padded slots: ball query pads slots >= cnt with the first hit (ball_query_gpu.cu:35-39), the
// MLP is row-wise and max-pool is idempotent under duplicated rows, so evaluating only the cnt genuine neighbours
// (or the single point 0 for an empty ball, whose slots are all 0) gives the same maximum as the reference's
// nsample rows.
//
// Work mapping: a warp owns 32 consecutive centres of one cloud.  Their genuine (centre, neighbour) rows are packed
// densely -- lane <-> row, 32 rows per pass -- so sparse neighbourhoods (LiDAR far field: 1-2 neighbours) do not idle
// 31 of 32 lanes.  After the MLP the 32x32 result tile goes through shared memory (stride 33, conflict free) so that
// lane <-> channel: the max over a centre's rows is a running maximum in a register, and each finished centre is one
// coalesced 128-byte store.  The 1.7 KB of weights per scale live in the kernel parameter block (constant bank), so an
// FFMA reads its weight as a constant operand instead of through a shared-memory load.
#include "common.cuh"
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "kernels.cuh"

namespace dclr {

constexpr int SA_WARPS = 4;
constexpr int SA_WMAX = 4 * 16 + 16 + 16 * 16 + 16 + 32 * 16 + 32;   // CIN <= 4, mlp [16,16,32]

struct SaParams {
    const float *x;          // [B,N,C] points (xyz + features), C = CIN
    const float *new_xyz;    // [B,M,3]
    const int32_t *idx[2];   // compact neighbour lists [B,M,ns]
    const int32_t *cnt[2];   // [B,M]
    int ns[2];
    float *out;              // [B,M,ld_out] point-major; scale s writes columns [s*32, (s+1)*32)
    uint16_t *out16;         // same layout as fp16 (fmt16 = 2) / bf16 (fmt16 = 1) for the tensor-core stages, or null
    int N, M, ld_out, nscales, fmt16;
    float w[2][SA_WMAX];     // per scale: W1[H1*CIN] b1[H1] W2[H2*H1] b2[H2] W3[H3*H2] b3[H3]
};

// CPW = centres per warp: 32 for full batches; 8 when the launch would otherwise leave most SMs empty (batch-1 latency:
// 2 clouds x 1024 centres are only 64 warps of 32 centres).  Lanes >= CPW own no centre; rows are packed the same way.
template <int CIN, int H1, int H2, int H3, int CPW>
__global__ void __launch_bounds__(SA_WARPS * 32)
sa_mlp_kernel(const __grid_constant__ SaParams a) {
    static_assert(H3 == 32, "lane <-> channel reduction assumes 32 output channels");
    __shared__ float s_tile[SA_WARPS][32][33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.y;
    const int c0 = (blockIdx.x * SA_WARPS + warp) * CPW;
    if (c0 >= a.M) return;
    const int c = c0 + lane;
    const bool has_c = lane < CPW && c < a.M;
    const float *ctr = a.new_xyz + ((size_t)b * a.M + (has_c ? c : c0)) * 3;
    const float ccx = __ldg(ctr), ccy = __ldg(ctr + 1), ccz = __ldg(ctr + 2);
    const float *xb = a.x + (size_t)b * a.N * CIN;
    float(*tile)[33] = s_tile[warp];

#pragma unroll
    for (int s = 0; s < 2; ++s) {
        if (s >= a.nscales) break;
        const float *W1 = a.w[s], *B1 = W1 + H1 * CIN, *W2 = B1 + H1, *B2 = W2 + H2 * H1, *W3 = B2 + H2,
                    *B3 = W3 + H3 * H2;
        const int n_raw = has_c ? __ldg(a.cnt[s] + (size_t)b * a.M + c) : 0;
        const int n_rows = has_c ? (n_raw > 0 ? n_raw : 1) : 0;       // empty ball: all slots are index 0
        int incl = n_rows;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        const int excl = incl - n_rows;
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        const int32_t *lists = a.idx[s] + ((size_t)b * a.M + c0) * a.ns[s];
        float cur = 0.f;                                             // post-ReLU values are >= 0
        int done = 0;                                                // centres of this warp already written
        for (int r0 = 0; r0 < total; r0 += 32) {
            const int row = r0 + lane;
            const bool act = row < total;
            // centre of this row: the last ci with excl[ci] <= row (binary search over the lanes' prefix sums)
            int ci = 0;
#pragma unroll
            for (int step = 16; step >= 1; step >>= 1) {
                const int probe = ci + step;
                const int e = __shfl_sync(0xffffffffu, excl, probe & 31);
                if (probe < 32 && e <= row) ci = probe;
            }
            // rows of lanes past a.M never exist (n_rows = 0 there), but equal prefix values must resolve to the
            // LAST centre with that prefix that owns rows: excl is non-decreasing and strictly increases after every
            // centre with rows, so "last ci with excl <= row" is the owner as long as row < total.
            const int j = row - __shfl_sync(0xffffffffu, excl, ci);
            const int raw_ci = __shfl_sync(0xffffffffu, n_raw, ci);
            const float cx = __shfl_sync(0xffffffffu, ccx, ci), cy = __shfl_sync(0xffffffffu, ccy, ci),
                        cz = __shfl_sync(0xffffffffu, ccz, ci);
            const int nrows_ci = __shfl_sync(0xffffffffu, n_rows, ci);
            const bool last_of_centre = act && (j == nrows_ci - 1);
            const unsigned lastmask = __ballot_sync(0xffffffffu, last_of_centre);
            const int k = (act && raw_ci > 0) ? __ldg(lists + (size_t)ci * a.ns[s] + j) : 0;
            float in[CIN];
            if (CIN == 4) {
                const float4 v = __ldg(reinterpret_cast<const float4 *>(xb) + k);
                in[0] = v.x; in[1] = v.y; in[2] = v.z; in[CIN - 1] = v.w;
            } else {
#pragma unroll
                for (int i = 0; i < CIN; ++i) in[i] = __ldg(xb + (size_t)k * CIN + i);
            }
            in[0] -= cx; in[1] -= cy; in[2] -= cz;                   // grouped_xyz -= new_xyz (pointnet2_utils.py:252)
            float h1[H1], h2[H2];
#pragma unroll
            for (int o = 0; o < H1; ++o) {
                float v = B1[o];
#pragma unroll
                for (int i = 0; i < CIN; ++i) v = fmaf(W1[o * CIN + i], in[i], v);
                h1[o] = fmaxf(v, 0.f);
            }
#pragma unroll
            for (int o = 0; o < H2; ++o) {
                float v = B2[o];
#pragma unroll
                for (int i = 0; i < H1; ++i) v = fmaf(W2[o * H1 + i], h1[i], v);
                h2[o] = fmaxf(v, 0.f);
            }
            __syncwarp();
#pragma unroll
            for (int o = 0; o < H3; ++o) {
                float v = B3[o];
#pragma unroll
                for (int i = 0; i < H2; ++i) v = fmaf(W3[o * H2 + i], h2[i], v);
                tile[lane][o] = fmaxf(v, 0.f);
            }
            __syncwarp();
            // lane <-> channel: running max over the rows, flushed at the last row of every centre
            const int nrow = min(32, total - r0);
            for (int r = 0; r < nrow; ++r) {
                cur = fmaxf(cur, tile[r][lane]);
                if ((lastmask >> r) & 1u) {
                    // centres without rows do not exist below a.M, so finished centres are consecutive
                    const size_t o = ((size_t)b * a.M + c0 + done) * a.ld_out + s * H3 + lane;
                    a.out[o] = cur;
                    if (a.out16)
                        a.out16[o] = a.fmt16 == 1 ? __bfloat16_as_ushort(__float2bfloat16_rn(cur))
                                                  : __half_as_ushort(__float2half_rn(cur));
                    cur = 0.f;
                    ++done;
                }
            }
        }
    }
}

int sa_mlp_launch(const SaArgs &s, int B, int cin, const int *h, cudaStream_t st) {
    if (!(h[0] == 16 && h[1] == 16 && h[2] == 32))
        return fail(DCLR_E_UNSUPPORTED, "sa_mlp: only mlp [16,16,32] is built (got [%d,%d,%d])", h[0], h[1], h[2]);
    if (cin != 3 && cin != 4) return fail(DCLR_E_UNSUPPORTED, "sa_mlp: input_dim must be 3 or 4 (got %d)", cin);
    SaParams p;
    p.x = s.x; p.new_xyz = s.new_xyz; p.out = s.out; p.out16 = s.out16; p.fmt16 = s.fmt16; p.N = s.N; p.M = s.M; p.ld_out = s.ld_out; p.nscales = s.nscales;
    const int wsz = cin * 16 + 16 + 16 * 16 + 16 + 32 * 16 + 32;
    for (int i = 0; i < 2; ++i) {
        p.idx[i] = s.idx[i]; p.cnt[i] = s.cnt[i]; p.ns[i] = s.ns[i];
        for (int e = 0; e < SA_WMAX; ++e) p.w[i][e] = (i < s.nscales && e < wsz) ? s.w_host[i][e] : 0.f;
    }
    // few centres in the whole launch (small batches): 8 centres per warp spread them over 4x as many warps
    const bool small = (long long)B * s.M < 148LL * SA_WARPS * 32;
    if (small) {
        dim3 grid(div_up(s.M, SA_WARPS * 8), B);
        if (cin == 4) sa_mlp_kernel<4, 16, 16, 32, 8><<<grid, SA_WARPS * 32, 0, st>>>(p);
        else sa_mlp_kernel<3, 16, 16, 32, 8><<<grid, SA_WARPS * 32, 0, st>>>(p);
    } else {
        dim3 grid(div_up(s.M, SA_WARPS * 32), B);
        if (cin == 4) sa_mlp_kernel<4, 16, 16, 32, 32><<<grid, SA_WARPS * 32, 0, st>>>(p);
        else sa_mlp_kernel<3, 16, 16, 32, 32><<<grid, SA_WARPS * 32, 0, st>>>(p);
    }
    DCLR_LAUNCH_CHECK("sa_mlp_kernel");
    return DCLR_OK;
}

}  // namespace dclr
