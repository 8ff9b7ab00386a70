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
// All activations that cross kernels are 16-bit images in the model's operand format (fp16 / bf16):
// feat16 [2B,n,F] point-major (templates first), nbr [B,n,k] row records {source - template displacement, source index}
// from knn_launch -> motion16 [B*n, merge_out] features only (the coordinates stay in xyz); motion32 (nullable)
// additionally receives the fp32 values for the API tap.
int tc_flow(TcModel *t, const uint16_t *feat16, const float4 *nbr, int B, int n, uint16_t *motion16, float *motion32,
            cudaStream_t st);
// motion16 [B*n, merge_out] + xyz [B*n,3] -> pooled [B, C_last] (max over the points); zero-fills pooled first
int tc_head(TcModel *t, const float *xyz, const uint16_t *motion16, int B, int n, float *pooled, cudaStream_t st);
// fp32 -> the model's 16-bit format, elementwise (features supplied by the caller: dclr_forward_features)
int tc_convert(TcModel *t, const float *src, uint16_t *dst, size_t n, cudaStream_t st);
int tc_format(const TcModel *t);

}  // namespace dclr
