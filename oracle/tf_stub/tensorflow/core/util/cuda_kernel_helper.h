// Stand-in for TensorFlow 1.x tensorflow/core/util/cuda_kernel_helper.h (third-party to the reference, not
// vendored there).  Restates the two helpers the reference kernels use:
//   CUDA_1D_KERNEL_LOOP(i, n)      grid-stride loop over [0, n)
//   GetCudaLaunchConfig(n, device) 1024-thread blocks, block count = min(ceil(min(SMs*maxThreadsPerSM, n)/1024), SMs)
// Launch shape does not change any result (the loops are grid-stride); it only matters for timing.
#pragma once
#include <cuda_runtime.h>
#include <algorithm>
#include "third_party/eigen3/unsupported/Eigen/CXX11/Tensor"
#define CUDA_1D_KERNEL_LOOP(i, n) \
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += blockDim.x * gridDim.x)
namespace tensorflow {
struct CudaLaunchConfig {
  int virtual_thread_count = -1;
  int thread_per_block = -1;
  int block_count = -1;
};
inline CudaLaunchConfig GetCudaLaunchConfig(int work_element_count, const Eigen::GpuDevice&) {
  int dev = 0, sms = 1, tpsm = 2048;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&tpsm, cudaDevAttrMaxThreadsPerMultiProcessor, dev);
  CudaLaunchConfig c;
  c.virtual_thread_count = work_element_count;
  const int physical = std::min(sms * tpsm, work_element_count);
  c.thread_per_block = 1024;
  c.block_count = std::min((physical + 1023) / 1024, sms);
  if (c.block_count < 1) c.block_count = 1;
  return c;
}
}  // namespace tensorflow
