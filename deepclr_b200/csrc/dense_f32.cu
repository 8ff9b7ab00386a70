// fp32 SIMT building blocks of the "fp32-faithful" numerics mode and the glue kernels shared by both modes:
// tiled Y = act(X W^T + b), flow-embedding row builder, masked max over neighbours, max over points, pose
// activation, layout transposes.  Reference semantics: deepclr/models/deepclr.py:201-231 (motion embedding),
// :284-294 (output), helper.py:11-38 (Conv1d k=1 + ReLU).
#include "common.cuh"
#include "kernels.cuh"

namespace dclr {

// ---------------------------------------------------------------- Y[M,N] = act(X[M,K] W[N,K]^T + b)
constexpr int LBM = 64, LBN = 64, LBK = 16;

__global__ void __launch_bounds__(256)
linear_f32_kernel(const float *__restrict__ X, int ldx, const float *__restrict__ W, const float *__restrict__ bias,
                  float *__restrict__ Y, int ldy, int M, int N, int K, int relu) {
    __shared__ float sX[LBK][LBM + 4];
    __shared__ float sW[LBK][LBN + 4];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;            // 16 x 16 threads, 4 x 4 outputs each
    const int m0 = blockIdx.y * LBM, n0 = blockIdx.x * LBN;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int k0 = 0; k0 < K; k0 += LBK) {
        // 64 rows x 16 k = 1024 elements per operand, 4 per thread; consecutive threads read consecutive k
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            int lin = tid + e * 256;
            int r = lin >> 4, kk = lin & 15;
            int gm = m0 + r, gn = n0 + r, gk = k0 + kk;
            sX[kk][r] = (gm < M && gk < K) ? __ldg(X + (size_t)gm * ldx + gk) : 0.f;
            sW[kk][r] = (gn < N && gk < K) ? __ldg(W + (size_t)gn * K + gk) : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < LBK; ++kk) {
            float xa[4], wb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) xa[i] = sX[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) wb[j] = sW[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(xa[i], wb[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int gm = m0 + ty * 4 + i;
        if (gm >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int gn = n0 + tx * 4 + j;
            if (gn >= N) continue;
            float v = acc[i][j] + __ldg(bias + gn);
            Y[(size_t)gm * ldy + gn] = relu ? fmaxf(v, 0.f) : v;
        }
    }
}

int linear_f32_launch(const float *X, int ldx, const float *W, const float *b, float *Y, int ldy, int M, int N, int K,
                      int relu, cudaStream_t st) {
    dim3 grid(div_up(N, LBN), div_up(M, LBM));
    linear_f32_kernel<<<grid, 256, 0, st>>>(X, ldx, W, b, Y, ldy, M, N, K, relu);
    DCLR_LAUNCH_CHECK("linear_f32_kernel");
    return DCLR_OK;
}

// ---------------------------------------------------------------- skinny linear for the pose tail (few rows)
// Y[M,N] = act(X[M,K] W[N,K]^T + b) when M is a batch of pooled vectors.  A warp owns SK_COLS output columns x SK_ROWS
// rows: lanes split K (coalesced float4 of W and X), 64 partial sums per lane, xor-shuffle trees at the end.  Every X
// value loaded feeds SK_COLS FMAs -- the first version (1 column x 32 rows per warp) streamed the whole X block
// through L1 once per output column (128 MB of L1 traffic for 64 x 1024 -> 512) and took ~19 us per layer.  The
// summation order of an output (k = 4 lane + 128 j ascending, then the shuffle tree) is the same as before.
// act: 0 none, 1 ReLU, 2 pose activation of `label_type` (deepclr.py:275-282).  Needs K % 4 == 0.
constexpr int SK_ROWS = 16, SK_COLS = 4, SK_WARPS = 4;

__device__ __forceinline__ float pose_act(float v, int c, int label_type) {
    if (label_type == 2) {
        if (c == 0) v = 1.f / (1.f + expf(-v));
        else if (c < 4) v = tanhf(v);
    } else if (label_type == 1) {
        if (c == 3) v = 1.f / (1.f + expf(-v));
        else if (c > 3) v = tanhf(v);
    }
    return v;
}

__global__ void __launch_bounds__(SK_WARPS * 32)
linear_skinny_kernel(const float *__restrict__ X, int ldx, const float *__restrict__ W, const float *__restrict__ bias,
                     float *__restrict__ Y, int ldy, int M, int N, int K, int act, int label_type) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n0 = (blockIdx.x * SK_WARPS + warp) * SK_COLS;
    const int m0 = blockIdx.y * SK_ROWS;
    if (n0 >= N) return;
    float acc[SK_COLS][SK_ROWS];
#pragma unroll
    for (int c = 0; c < SK_COLS; ++c)
#pragma unroll
        for (int i = 0; i < SK_ROWS; ++i) acc[c][i] = 0.f;
    const float *wrow[SK_COLS];
#pragma unroll
    for (int c = 0; c < SK_COLS; ++c) wrow[c] = W + (size_t)(n0 + c < N ? n0 + c : N - 1) * K;
    for (int k0 = lane * 4; k0 < K; k0 += 128) {
        float4 w[SK_COLS];
#pragma unroll
        for (int c = 0; c < SK_COLS; ++c) w[c] = __ldg(reinterpret_cast<const float4 *>(wrow[c] + k0));
#pragma unroll
        for (int i = 0; i < SK_ROWS; ++i) {
            const int m = m0 + i < M ? m0 + i : M - 1;
            const float4 xv = __ldg(reinterpret_cast<const float4 *>(X + (size_t)m * ldx + k0));
#pragma unroll
            for (int c = 0; c < SK_COLS; ++c) {
                acc[c][i] = fmaf(w[c].x, xv.x, acc[c][i]);
                acc[c][i] = fmaf(w[c].y, xv.y, acc[c][i]);
                acc[c][i] = fmaf(w[c].z, xv.z, acc[c][i]);
                acc[c][i] = fmaf(w[c].w, xv.w, acc[c][i]);
            }
        }
    }
    // output (c, i) lands in lane (c SK_ROWS + i) % 32, slot (c SK_ROWS + i) / 32
    float mine[SK_COLS * SK_ROWS / 32];
#pragma unroll
    for (int o = 0; o < SK_COLS * SK_ROWS / 32; ++o) mine[o] = 0.f;
#pragma unroll
    for (int c = 0; c < SK_COLS; ++c)
#pragma unroll
        for (int i = 0; i < SK_ROWS; ++i) {
            float v = acc[c][i];
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
            if (lane == (c * SK_ROWS + i) % 32) mine[(c * SK_ROWS + i) / 32] = v;
        }
#pragma unroll
    for (int o = 0; o < SK_COLS * SK_ROWS / 32; ++o) {
        const int e = o * 32 + lane, c = e / SK_ROWS, i = e % SK_ROWS;
        const int m = m0 + i, n = n0 + c;
        if (m < M && n < N) {
            float v = mine[o] + __ldg(bias + n);
            if (act == 1) v = fmaxf(v, 0.f);
            else if (act == 2) v = pose_act(v, n, label_type);
            Y[(size_t)m * ldy + n] = v;
        }
    }
}

int linear_skinny_launch(const float *X, int ldx, const float *W, const float *b, float *Y, int ldy, int M, int N,
                         int K, int act, int label_type, cudaStream_t st) {
    if (K % 4 || ldx % 4) return fail(DCLR_E_UNSUPPORTED, "linear_skinny: K and ldx must be multiples of 4");
    dim3 grid(div_up(N, SK_WARPS * SK_COLS), div_up(M, SK_ROWS));
    linear_skinny_kernel<<<grid, SK_WARPS * 32, 0, st>>>(X, ldx, W, b, Y, ldy, M, N, K, act, label_type);
    DCLR_LAUNCH_CHECK("linear_skinny_kernel");
    return DCLR_OK;
}

// ---------------------------------------------------------------- flow embedding rows
// row (b,i,j): [ p_src(nbr) - p_tmpl(i) | f_tmpl(i) | f_src(nbr) ]  (deepclr.py:206-210), plus the radius mask
// (deepclr.py:220-222).  xyz [2B,n,3], feat [2B,n,F] point-major, templates first.
__global__ void flow_rows_kernel(const float *__restrict__ xyz, const float *__restrict__ feat,
                                 const int32_t *__restrict__ nbr, int B, int n, int k, int F, float radius,
                                 float *__restrict__ rows, int ld, uint8_t *__restrict__ mask) {
    const size_t row = blockIdx.x;                 // (b*n + i)*k + j
    const int j = (int)(row % k);
    const size_t bi = row / k;
    const int b = (int)(bi / n);
    const int src = __ldg(nbr + row);
    (void)j;
    const float *pt = xyz + bi * 3;
    const float *ps = xyz + ((size_t)(B + b) * n + src) * 3;
    const float *ft = feat + bi * F;
    const float *fs = feat + ((size_t)(B + b) * n + src) * F;
    float *out = rows + row * ld;
    const int D = 3 + 2 * F;
    for (int c = threadIdx.x; c < ld; c += blockDim.x) {
        float v = 0.f;
        if (c < 3) v = __ldg(ps + c) - __ldg(pt + c);
        else if (c < 3 + F) v = __ldg(ft + c - 3);
        else if (c < D) v = __ldg(fs + c - 3 - F);
        out[c] = v;
    }
    if (threadIdx.x == 0) {
        float dx = __ldg(ps) - __ldg(pt), dy = __ldg(ps + 1) - __ldg(pt + 1), dz = __ldg(ps + 2) - __ldg(pt + 2);
        float nrm = sqrtf(dx * dx + dy * dy + dz * dz);
        mask[row] = (radius > 0.f && nrm >= radius) ? 1 : 0;
    }
}

int flow_rows_launch(const float *xyz, const float *feat, const int32_t *nbr, int B, int n, int k, int F, float radius,
                     float *rows, int ld, uint8_t *mask, cudaStream_t st) {
    flow_rows_kernel<<<(unsigned)((size_t)B * n * k), 64, 0, st>>>(xyz, feat, nbr, B, n, k, F, radius, rows, ld, mask);
    DCLR_LAUNCH_CHECK("flow_rows_kernel");
    return DCLR_OK;
}

// out[g, 0:3] = xyz_tmpl[g]; out[g, 3+c] = max_j (mask ? 0 : H[(g*k+j), c])   (deepclr.py:220-229)
__global__ void masked_groupmax_kernel(const float *__restrict__ H, int ldh, const uint8_t *__restrict__ mask,
                                       const float *__restrict__ xyz, int k, int C, float *__restrict__ out, int ldo) {
    const size_t g = blockIdx.x;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float m = 0.f;     // k >= 1 and values are post-ReLU, masked entries are exactly 0
        for (int j = 0; j < k; ++j) {
            float v = __ldg(H + (g * k + j) * ldh + c);
            m = fmaxf(m, mask[g * k + j] ? 0.f : v);
        }
        out[g * ldo + 3 + c] = m;
    }
    if (threadIdx.x < 3) out[g * ldo + threadIdx.x] = __ldg(xyz + g * 3 + threadIdx.x);
    for (int c = 3 + C + threadIdx.x; c < ldo; c += blockDim.x) out[g * ldo + c] = 0.f;
}

int masked_groupmax_launch(const float *H, int ldh, const uint8_t *mask, const float *xyz, size_t groups, int k, int C,
                           float *out, int ldo, cudaStream_t st) {
    masked_groupmax_kernel<<<(unsigned)groups, 128, 0, st>>>(H, ldh, mask, xyz, k, C, out, ldo);
    DCLR_LAUNCH_CHECK("masked_groupmax_kernel");
    return DCLR_OK;
}

// out[b,c] = max_i H[b*n+i, c]   (deepclr.py:287)
__global__ void rowmax_kernel(const float *__restrict__ H, int ldh, int n, int C, float *__restrict__ out) {
    __shared__ float s[8][33];
    const int b = blockIdx.y;
    const int c = blockIdx.x * 32 + (threadIdx.x & 31);
    const int w = threadIdx.x >> 5;
    float m = -INFINITY;
    if (c < C)
        for (int i = w; i < n; i += 8) m = fmaxf(m, __ldg(H + ((size_t)b * n + i) * ldh + c));
    s[w][threadIdx.x & 31] = m;
    __syncthreads();
    if (w == 0 && c < C) {
#pragma unroll
        for (int q = 1; q < 8; ++q) m = fmaxf(m, s[q][threadIdx.x]);
        out[(size_t)b * C + c] = m;
    }
}

int rowmax_launch(const float *H, int ldh, int B, int n, int C, float *out, cudaStream_t st) {
    rowmax_kernel<<<dim3(div_up(C, 32), B), 256, 0, st>>>(H, ldh, n, C, out);
    DCLR_LAUNCH_CHECK("rowmax_kernel");
    return DCLR_OK;
}

// deepclr.py:275-282
__global__ void pose_activation_kernel(float *__restrict__ y, int B, int dim, int label_type) {
    int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= B * dim) return;
    int c = e % dim;
    float v = y[e];
    if (label_type == 2) {
        if (c == 0) v = 1.f / (1.f + expf(-v));
        else if (c < 4) v = tanhf(v);
    } else if (label_type == 1) {
        if (c == 3) v = 1.f / (1.f + expf(-v));
        else if (c > 3) v = tanhf(v);
    }
    y[e] = v;
}

int pose_activation_launch(float *y, int B, int dim, int label_type, cudaStream_t st) {
    pose_activation_kernel<<<div_up(B * dim, 128), 128, 0, st>>>(y, B, dim, label_type);
    DCLR_LAUNCH_CHECK("pose_activation_kernel");
    return DCLR_OK;
}

// ---------------------------------------------------------------- layout glue
// point-major (xyz [G,n,3], feat [G,n,ldf] first F cols) -> channel-major out [G, 3+F, n]   (deepclr.py:40-45)
__global__ void pm_to_cm_kernel(const float *__restrict__ xyz, const float *__restrict__ feat, int ldf, int off, int n,
                                int F, float *__restrict__ out) {
    __shared__ float tile[32][33];
    const int g = blockIdx.z;
    const int i0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int C = 3 + F;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        int i = i0 + r, c = c0 + threadIdx.x;
        float v = 0.f;
        if (i < n && c < C) {
            if (xyz) v = c < 3 ? __ldg(xyz + ((size_t)g * n + i) * 3 + c) : __ldg(feat + ((size_t)g * n + i) * ldf + off + c - 3);
            else v = __ldg(feat + ((size_t)g * n + i) * ldf + c);
        }
        tile[r][threadIdx.x] = v;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        int c = c0 + r, i = i0 + threadIdx.x;
        if (c < C && i < n) out[((size_t)g * C + c) * n + i] = tile[threadIdx.x][r];
    }
}

int pm_to_cm_launch(const float *xyz, const float *feat, int ldf, int off, int G, int n, int F, float *out,
                    cudaStream_t st) {
    dim3 grid(div_up(n, 32), div_up(3 + F, 32), G);
    pm_to_cm_kernel<<<grid, dim3(32, 8), 0, st>>>(xyz, feat, ldf, off, n, F, out);
    DCLR_LAUNCH_CHECK("pm_to_cm_kernel");
    return DCLR_OK;
}

// channel-major in [G, 3+F, n] -> xyz [G,n,3], feat [G,n,ldf] (F columns); xyz == nullptr: rows [xyz | features] in feat
__global__ void cm_to_pm_kernel(const float *__restrict__ in, int n, int F, float *__restrict__ xyz,
                                float *__restrict__ feat, int ldf) {
    __shared__ float tile[32][33];
    const int g = blockIdx.z;
    const int i0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int C = 3 + F;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        int c = c0 + r, i = i0 + threadIdx.x;
        tile[r][threadIdx.x] = (c < C && i < n) ? __ldg(in + ((size_t)g * C + c) * n + i) : 0.f;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        int i = i0 + r, c = c0 + threadIdx.x;
        if (i < n && c < C) {
            float v = tile[threadIdx.x][r];
            if (!xyz) feat[((size_t)g * n + i) * ldf + c] = v;
            else if (c < 3) xyz[((size_t)g * n + i) * 3 + c] = v;
            else feat[((size_t)g * n + i) * ldf + c - 3] = v;
        }
    }
}

int cm_to_pm_launch(const float *in, int G, int n, int F, float *xyz, float *feat, int ldf, cudaStream_t st) {
    dim3 grid(div_up(n, 32), div_up(3 + F, 32), G);
    cm_to_pm_kernel<<<grid, dim3(32, 8), 0, st>>>(in, n, F, xyz, feat, ldf);
    DCLR_LAUNCH_CHECK("cm_to_pm_kernel");
    return DCLR_OK;
}

}  // namespace dclr
