// corr_simt.cu -- tiled fp32 CUDA-core correlation for kernel_size == 1, stride_1 == 1 (placeholder: not yet enabled)
#include "fn2_common.cuh"
namespace fn2 {
bool corr_simt_supported(const CorrGeom&) { return false; }
int corr_simt_forward(const CorrGeom&, const float*, const float*, float*, cudaStream_t) { return FN2_E_UNSUPPORTED_PRECISION; }
int corr_simt_backward(const CorrGeom&, const float*, const float*, const float*, float*, float*, cudaStream_t) { return FN2_E_UNSUPPORTED_PRECISION; }
}  // namespace fn2
