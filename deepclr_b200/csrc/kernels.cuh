// Internal launch interfaces between the translation units of libdeepclr_b200.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

namespace dclr {

// fps.cu
// tie_rule: DCLR_TIE_LOWEST_INDEX / DCLR_TIE_REFERENCE (include/deepclr_b200.h); an argument of every launch, no global state
int fps_launch(const float *pts, int stride, int B, int N, int npoint, int tie_rule, int32_t *idx, float *new_xyz,
               float *temp_big, cudaStream_t st);

// fps_bucket.cu: exact pruned FPS for 8192 < N <= 131072 (smaller clouds: the resident kernel is faster); *used = 0
// when the shape is not handled
size_t fps_bucket_workspace_bytes(int B, int N);
int fps_bucket_l2_launch(const float4 *sorted, float *tmin, int B, int N, int npoint, int tieL, int32_t *idx, float *nx,
                         cudaStream_t st);
// fps_tmem.cu: running minima in tensor memory, up to mp samples per round; sorted = fps_sort_kernel's records, N <= 16384
int fps_tmem_launch(const float4 *sorted, int B, int N, int npoint, int tie_ref, int nw, int mp, int32_t *idx, float *nx,
                    cudaStream_t st);
// fps_cluster.cu: the same, one cloud split over a cluster of two CTAs (8192 < N <= 16384)
int fps_cluster_launch(const float4 *sorted, int B, int N, int npoint, int tie_ref, int mp, int32_t *idx, float *nx,
                       cudaStream_t st);
// after_sort (optional) is recorded on st between the sort and the sampling kernel
int fps_bucket_launch(const float *pts, int stride, int B, int N, int npoint, int tie_rule, int32_t *idx, float *new_xyz,
                      void *ws, cudaStream_t st, int *used, cudaEvent_t after_sort = nullptr);

// pointops.cu
int ball_query_launch(const float *new_xyz, const float *pts, int stride, int B, int N, int M, int nr,
                      const float *radii, const int *ns, int32_t *const *idx, int32_t *const *cnt, int pad,
                      cudaStream_t st);
// ballgrid.cu: grid-accelerated ball query for large clouds; *used = 0 when the shape is not handled (then call
// ball_query_launch).  ws must hold ball_grid_workspace_bytes(B, N) bytes.
constexpr int BG_MIN_N = 1024;          // smallest cloud that takes the uniform grid (ballgrid.cu)
size_t ball_grid_workspace_bytes(int B, int N);
// phase: the grid build reads only the cloud, so the model runs it on a side stream next to FPS (BG_PHASE_BUILD, new_xyz
// unused) and the query after both (BG_PHASE_QUERY); *used is the same for every phase of a shape.
enum { BG_PHASE_BOTH = 0, BG_PHASE_BUILD = 1, BG_PHASE_QUERY = 2 };
int ball_grid_launch(const float *new_xyz, const float *pts, int stride, int B, int N, int M, int nr,
                     const float *radii, const int *ns, int32_t *const *idx, int32_t *const *cnt, int pad, void *ws,
                     cudaStream_t st, int *used, int phase = BG_PHASE_BOTH);
// nbr4 (optional) [B,Ny,k]: {x_cand - y_query (3 floats, the subtraction of the distance), candidate index bits} --
// what the tensor-core flow kernel needs per row, in one aligned 16-byte record
int knn_launch(const float *x, const float *y, int B, int Nx, int Ny, int k, int64_t *row, int64_t *col,
               int32_t *col_local, float4 *nbr4, cudaStream_t st);

// sa_mlp.cu
struct SaArgs {
    const float *x;          // [B,N,C] points (xyz + features), C = CIN
    const float *new_xyz;    // [B,M,3]
    const int32_t *idx[2];   // compact neighbour lists [B,M,ns]
    const int32_t *cnt[2];   // [B,M]
    int ns[2];
    const float *w_host[2];  // HOST copies, per scale: W1[H1*CIN] b1[H1] W2[H2*H1] b2[H2] W3[H3*H2] b3[H3]
                             // (they travel in the kernel parameter block = constant bank)
    float *out;              // [B,M,ld_out] point-major; scale s writes columns [s*H3, (s+1)*H3)
    uint16_t *out16 = nullptr;   // optional 16-bit copy (fmt16: 1 bf16, 2 fp16) read by the tensor-core flow kernel
    int fmt16 = 0;
    int N, M, ld_out, nscales;
};
int sa_mlp_launch(const SaArgs &a, int B, int cin, const int *h, cudaStream_t st);
// sa_tc.cu: the same stage on the tcgen05 tensor cores (fp16 hi/lo split operands, fp32-class results) for dense
// neighbourhoods; w_dev = 2 x sa_tc_weight_bytes() / 2 packed by sa_tc_pack_scale (host) per scale
size_t sa_tc_weight_bytes();
void sa_tc_pack_scale(const float *w_host, int cin, uint16_t *img);
int sa_tc_launch(const SaArgs &a, const uint16_t *w_dev, int B, int cin, cudaStream_t st);
int sa_count_rows_launch(const int32_t *cnt, size_t n, unsigned long long *rows, cudaStream_t st);

// dense_f32.cu
int linear_f32_launch(const float *X, int ldx, const float *W, const float *b, float *Y, int ldy, int M, int N, int K,
                      int relu, cudaStream_t st);
int linear_skinny_launch(const float *X, int ldx, const float *W, const float *b, float *Y, int ldy, int M, int N,
                         int K, int act, int label_type, cudaStream_t st);
int flow_rows_launch(const float *xyz, const float *feat, const int32_t *nbr, int B, int n, int k, int F, float radius,
                     float *rows, int ld, uint8_t *mask, cudaStream_t st);
int masked_groupmax_launch(const float *H, int ldh, const uint8_t *mask, const float *xyz, size_t groups, int k, int C,
                           float *out, int ldo, cudaStream_t st);
int rowmax_launch(const float *H, int ldh, int B, int n, int C, float *out, cudaStream_t st);
int pose_activation_launch(float *y, int B, int dim, int label_type, cudaStream_t st);
int pm_to_cm_launch(const float *xyz, const float *feat, int ldf, int off, int G, int n, int F, float *out,
                    cudaStream_t st);
int cm_to_pm_launch(const float *in, int G, int n, int F, float *xyz, float *feat, int ldf, cudaStream_t st);

}  // namespace dclr
