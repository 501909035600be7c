// fn2_downsample.cuh -- coordinate and weight arithmetic shared by the downsample kernels (downsample.cu,
// downsample_multi.cu).  Every function restates one expression of downsample_kernel_gpu.cu.cc:35-76.
#pragma once
#include "fn2_common.cuh"

namespace fn2 {

__device__ __forceinline__ float nan_f() { return __int_as_float(0x7fffffff); }   // :74

__device__ __forceinline__ float src_pos(int d, int on1, int in1) {
  // downsample_kernel_gpu.cu.cc:40-41, this evaluation order; __fmul_rn keeps the product a rounded fp32 value (the
  // compiler would otherwise fuse it into the subtraction of tri() and shift every weight by up to 4e-6)
  return __fmul_rn(__fdiv_rn((float)d, (float)on1), (float)in1);
}
__device__ __forceinline__ float tri(float off, float src, float scale) {
  return fmaxf(0.0f, 1.0f - (fabsf(off - src) / scale));   // :59-60
}
// last output index d in [0, on) whose centre is <= pos (centres are increasing); 0 if none
__device__ __forceinline__ int last_centre_le(int pos, float scale, int on, int in_) {
  int d = min(on - 1, max(0, (int)((float)pos / scale)));
  while (d + 1 < on && src_pos(d + 1, on - 1, in_ - 1) <= (float)pos) ++d;
  while (d > 0 && src_pos(d, on - 1, in_ - 1) > (float)pos) --d;
  return d;
}

}  // namespace fn2
