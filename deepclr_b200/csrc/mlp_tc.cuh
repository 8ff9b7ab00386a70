// bf16 tcgen05/TMEM path for the motion-embedding and output-head MLP stacks (mlp_tc.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dclr {

struct TcLayerSpec { const float *W, *b; int out, in; };   // device fp32 parameters, W row-major [out,in]
struct TcModel;

// specs[0..2] = merge MLP, specs[3..7] = output conv MLP
// fmt: 1 = bf16, 2 = fp16 operands (fp32 accumulation either way)
int tc_create(const TcLayerSpec *specs, int F, int k, float radius, int fmt, TcModel **out);
void tc_destroy(TcModel *t);
// xyz [2B,n,3], feat [2B,n,F] point-major (templates first), nbr [B,n,k] -> motion [B*n, merge_out] features only
// (the coordinates stay in xyz)
int tc_flow(TcModel *t, const float *xyz, const float *feat, const int32_t *nbr, int B, int n, float *motion,
            cudaStream_t st);
// motion [B*n, merge_out] + xyz [B*n,3] -> pooled [B, C_last] (max over the points); zero-fills pooled first
int tc_head(TcModel *t, const float *xyz, const float *motion, int B, int n, float *pooled, cudaStream_t st);
int tc_format(const TcModel *t);

}  // namespace dclr
