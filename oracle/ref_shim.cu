// ref_shim.cu -- C entry points around the reference's own CUDA launchers.  TEST INFRASTRUCTURE ONLY.
//
// oracle/Makefile compiles this file together with the reference's kernel sources, unmodified and where they lie
// under /root/reference/src/ops, into oracle/_ref/libfn2ref.so.  The shim restates the host side of each
// OpKernel::Compute (allocate padded temporaries, Pad x2, launch), because those need the TensorFlow runtime:
//   correlation fwd  src/ops/correlation/correlation_kernel.cc:35-123
//   correlation bwd  src/ops/correlation/correlation_grad_kernel.cc:33-146
//   flow_warp        src/ops/flow_warp/flow_warp.cc:17-42, flow_warp_grad.cc:17-51
//   downsample       src/ops/downsample/downsample_kernel.cc:22-41
// Everything runs on the legacy default stream, which is also where the reference's stray cudaMemset calls go.
#include <cuda_runtime.h>
#include <cmath>
#include "correlation_kernel.h"
#include "pad.h"
#include "flow_warp.h"
#include "downsample_kernel.h"

using tensorflow::GPUDevice;

extern "C" {

int fn2ref_correlation_forward(const float* a, const float* b, float* out, int N, int H, int W, int C, int k, int md,
                               int s1, int s2, int pad, float* pad_a, float* pad_b) {
  GPUDevice dev;
  int pH = H + 2 * pad, pW = W + 2 * pad, kr = (k - 1) / 2, border = md + kr;
  int oH = (int)ceil((float)(pH - border * 2) / (float)s1), oW = (int)ceil((float)(pW - border * 2) / (float)s1);
  if (oH < 1 || oW < 1) return -1;
  int r = md / s2, gw = r * 2 + 1, D = gw * gw;
  tensorflow::Pad(dev, a, N, H, W, C, pH, pW, pad_a);
  tensorflow::Pad(dev, b, N, H, W, C, pH, pW, pad_b);
  tensorflow::Correlation(dev, pad_a, pad_b, N, oH, oW, D, oH * oW * D, pH, pW, C, md, r, gw, kr, k, s1, s2, out);
  return cudaGetLastError() == cudaSuccess ? 0 : -100;
}

int fn2ref_correlation_backward(const float* g, const float* a, const float* b, float* ga, float* gb, int N, int H,
                                int W, int C, int k, int md, int s1, int s2, int pad, float* pad_a, float* pad_b) {
  GPUDevice dev;
  int pH = H + 2 * pad, pW = W + 2 * pad, kr = (k - 1) / 2, border = md + kr;
  int oH = (int)ceil((float)(pH - border * 2) / (float)s1), oW = (int)ceil((float)(pW - border * 2) / (float)s1);
  if (oH < 1 || oW < 1) return -1;
  int r = md / s2, gw = r * 2 + 1, D = gw * gw;
  tensorflow::Pad(dev, a, N, H, W, C, pH, pW, pad_a);
  tensorflow::Pad(dev, b, N, H, W, C, pH, pW, pad_b);
  tensorflow::CorrelationGradA(dev, N, oW, oH, D, md, r, gw, kr, s1, s2, W, H, pW, pH, C, H * W * C, pad, pad_b, g, ga);
  tensorflow::CorrelationGradB(dev, N, oW, oH, D, md, r, gw, kr, s1, s2, W, H, pW, pH, C, H * W * C, pad, pad_a, g, gb);
  return cudaGetLastError() == cudaSuccess ? 0 : -100;
}

int fn2ref_flow_warp_forward(const float* image, const float* flow, float* out, int N, int H, int W, int C) {
  GPUDevice dev;
  tensorflow::TTypes<float, 4>::ConstTensor ti{image, {N, H, W, C}}, tf{flow, {N, H, W, 2}};
  tensorflow::TTypes<float, 4>::Tensor to{out, {N, H, W, C}};
  tensorflow::FlowWarp(dev, ti, tf, to);
  return cudaGetLastError() == cudaSuccess ? 0 : -100;
}

int fn2ref_flow_warp_backward(const float* image, const float* flow, const float* g, float* gimage, float* gflow,
                              int N, int H, int W, int C) {
  GPUDevice dev;
  tensorflow::TTypes<float, 4>::ConstTensor ti{image, {N, H, W, C}}, tf{flow, {N, H, W, 2}}, tg{g, {N, H, W, C}};
  tensorflow::TTypes<float, 4>::Tensor gi{gimage, {N, H, W, C}}, gf{gflow, {N, H, W, 2}};
  tensorflow::FlowWarpGrad(dev, ti, tf, tg, gi, gf);
  return cudaGetLastError() == cudaSuccess ? 0 : -100;
}

int fn2ref_downsample_forward(const float* in, float* out, int N, int iH, int iW, int C, int oH, int oW) {
  GPUDevice dev;
  tensorflow::TTypes<float, 4>::ConstTensor ti{in, {N, iH, iW, C}};
  tensorflow::TTypes<float, 4>::Tensor to{out, {N, oH, oW, C}};
  return tensorflow::Downsample(dev, ti, to) ? 0 : -100;
}

int fn2ref_version(void) { return 1; }
}
