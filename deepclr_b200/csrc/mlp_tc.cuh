// bf16 tcgen05/TMEM path for the motion-embedding and output-head MLP stacks (mlp_tc.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dclr {

struct TcLayerSpec { const float *W, *b; int out, in; };   // device fp32 parameters, W row-major [out,in]
struct TcModel;

// specs[0..2] = merge MLP, specs[3..7] = output conv MLP
int tc_create(const TcLayerSpec *specs, int F, int k, float radius, TcModel **out);
void tc_destroy(TcModel *t);
// xyz [2B,n,3], feat [2B,n,F] point-major, nbr [B,n,k] -> motion [B*n, ldm] (xyz | features) and pooled [B, C_last]
int tc_merge_output(TcModel *t, const float *xyz, const float *feat, const int32_t *nbr, int B, int n, float *motion,
                    int ldm, float *pooled, cudaStream_t st, int *launches);

}  // namespace dclr
