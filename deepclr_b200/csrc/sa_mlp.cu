// Set-abstraction grouping + shared MLP + max-pool, fused (fp32 SIMT): one warp per FPS centre, one lane per
// neighbour.  Replaces, per scale, grouping_operation x2 + cat + 3x(Conv2d 1x1 + bias + ReLU) + max_pool2d of the
// reference (pointnet2_utils.py:249-257, pytorch_utils.py:21-32, pointnet2_modules.py:40-52): the grouped tensor
// [B, 3+C, npoint, nsample] is never formed.
//
// Exactness of skipping padded slots: ball query pads slots >= cnt with the first hit (ball_query_gpu.cu:35-39), the
// MLP is row-wise and max-pool is idempotent under duplicated rows, so evaluating only the cnt genuine neighbours
// (or the single point 0 for an empty ball, whose slots are all 0) gives the same maximum as the reference's
// nsample rows.
#include "common.cuh"
#include "kernels.cuh"

namespace dclr {

constexpr int SA_WARPS = 8;

template <int CIN, int H1, int H2, int H3>
__global__ void __launch_bounds__(SA_WARPS * 32)
sa_mlp_kernel(SaArgs a) {
    static_assert(H3 == 32, "warp transpose-reduce below assumes 32 output channels");
    constexpr int WSZ = H1 * CIN + H1 + H2 * H1 + H2 + H3 * H2 + H3;
    __shared__ __align__(16) float s_w[2][WSZ];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.y;
    for (int s = 0; s < a.nscales; ++s)
        for (int e = tid; e < WSZ; e += SA_WARPS * 32) s_w[s][e] = __ldg(a.w[s] + e);
    __syncthreads();
    const int c = blockIdx.x * SA_WARPS + warp;
    if (c >= a.M) return;
    const float *ctr = a.new_xyz + ((size_t)b * a.M + c) * 3;
    const float cx = __ldg(ctr), cy = __ldg(ctr + 1), cz = __ldg(ctr + 2);
    const float *xb = a.x + (size_t)b * a.N * CIN;

    for (int s = 0; s < a.nscales; ++s) {
        const float *W1 = s_w[s], *B1 = W1 + H1 * CIN, *W2 = B1 + H1, *B2 = W2 + H2 * H1, *W3 = B2 + H2,
                    *B3 = W3 + H3 * H2;
        const int n_raw = __ldg(a.cnt[s] + (size_t)b * a.M + c);
        const int n = n_raw > 0 ? n_raw : 1;               // empty ball: all slots are index 0
        const int32_t *list = a.idx[s] + ((size_t)b * a.M + c) * a.ns[s];
        float acc[H3];
#pragma unroll
        for (int o = 0; o < H3; ++o) acc[o] = 0.f;          // post-ReLU values are >= 0
        for (int j0 = 0; j0 < n; j0 += 32) {
            const int j = j0 + lane;
            const bool act = j < n;
            const int k = (act && n_raw > 0) ? __ldg(list + j) : 0;
            float in[CIN];
            if (CIN == 4) {
                float4 v = __ldg(reinterpret_cast<const float4 *>(xb) + k);
                in[0] = v.x; in[1] = v.y; in[2] = v.z; in[CIN - 1] = v.w;
            } else {
#pragma unroll
                for (int i = 0; i < CIN; ++i) in[i] = __ldg(xb + (size_t)k * CIN + i);
            }
            in[0] -= cx; in[1] -= cy; in[2] -= cz;          // grouped_xyz -= new_xyz (pointnet2_utils.py:252)
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
#pragma unroll
            for (int o = 0; o < H3; ++o) {
                float v = B3[o];
#pragma unroll
                for (int i = 0; i < H2; ++i) v = fmaf(W3[o * H2 + i], h2[i], v);
                v = fmaxf(v, 0.f);
                acc[o] = act ? fmaxf(acc[o], v) : acc[o];
            }
        }
        // transpose-reduce: after the five steps lane L holds max over lanes of channel L in acc[0]
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
            const bool up = (lane & d) != 0;
#pragma unroll
            for (int i = 0; i < d; ++i) {
                const float send = up ? acc[i] : acc[i + d];
                const float keep = up ? acc[i + d] : acc[i];
                acc[i] = fmaxf(keep, __shfl_xor_sync(0xffffffffu, send, d));
            }
        }
        a.out[((size_t)b * a.M + c) * a.ld_out + s * H3 + lane] = acc[0];
    }
}

int sa_mlp_launch(const SaArgs &a, int B, int cin, const int *h, cudaStream_t st) {
    if (!(h[0] == 16 && h[1] == 16 && h[2] == 32))
        return fail(DCLR_E_UNSUPPORTED, "sa_mlp: only mlp [16,16,32] is built (got [%d,%d,%d])", h[0], h[1], h[2]);
    dim3 grid(div_up(a.M, SA_WARPS), B);
    if (cin == 4) sa_mlp_kernel<4, 16, 16, 32><<<grid, SA_WARPS * 32, 0, st>>>(a);
    else if (cin == 3) sa_mlp_kernel<3, 16, 16, 32><<<grid, SA_WARPS * 32, 0, st>>>(a);
    else return fail(DCLR_E_UNSUPPORTED, "sa_mlp: input_dim must be 3 or 4 (got %d)", cin);
    DCLR_LAUNCH_CHECK("sa_mlp_kernel");
    return DCLR_OK;
}

}  // namespace dclr
