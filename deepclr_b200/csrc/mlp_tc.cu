// placeholder until the tcgen05 kernels land
#include "common.cuh"
#include "mlp_tc.cuh"
namespace dclr {
int tc_create(const TcLayerSpec *, int, int, float, TcModel **) {
    return fail(DCLR_E_UNSUPPORTED, "tensor-core path not built yet");
}
void tc_destroy(TcModel *) {}
int tc_merge_output(TcModel *, const float *, const float *, const int32_t *, int, int, float *, int, float *,
                    cudaStream_t, int *) {
    return fail(DCLR_E_UNSUPPORTED, "tensor-core path not built yet");
}
}  // namespace dclr
