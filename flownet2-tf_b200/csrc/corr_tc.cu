// corr_tc.cu -- tcgen05 correlation (placeholder: not yet enabled)
#include "fn2_common.cuh"
namespace fn2 {
bool corr_tc_supported(const CorrGeom&, int) { return false; }
size_t corr_tc_workspace_bytes(const CorrGeom&, int, int) { return 0; }
int corr_tc_forward(const CorrGeom&, const float*, const float*, float*, int, void*, size_t, cudaStream_t) { return FN2_E_UNSUPPORTED_PRECISION; }
int corr_tc_backward(const CorrGeom&, const float*, const float*, const float*, float*, float*, int, void*, size_t, cudaStream_t) { return FN2_E_UNSUPPORTED_PRECISION; }
}  // namespace fn2
