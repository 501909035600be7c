// Stand-in: TTypes<T,N>::Tensor / ConstTensor with data() and dimension(i), as used by
// src/ops/flow_warp/flow_warp.cu.cc:103-106 and src/ops/downsample/downsample_kernel_gpu.cu.cc:83-90.
#pragma once
namespace tensorflow {
template <typename T, int NDIMS>
struct FnTensorView {
  T* ptr;
  long dims[NDIMS];
  T* data() const { return ptr; }
  long dimension(int i) const { return dims[i]; }
};
template <typename T, int NDIMS = 1>
struct TTypes {
  typedef FnTensorView<T, NDIMS> Tensor;
  typedef FnTensorView<const T, NDIMS> ConstTensor;
};
}  // namespace tensorflow
