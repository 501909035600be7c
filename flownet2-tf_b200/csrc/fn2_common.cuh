// fn2_common.cuh -- shared declarations for libfn2ops.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <atomic>
#include "../../include/fn2ops.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libfn2ops is written for sm_100a (B200) only"
#endif

namespace fn2 {

// Geometry of one correlation call, derived exactly as correlation_kernel.cc:35-57 does.
struct CorrGeom {
  int N, H, W, C;
  int k, md, s1, s2, pad;
  int kr, oH, oW, r, gw, D;
  int off;  // unpadded coordinate of output (0,0)'s patch corner: md - pad
  // Row halos (fn2_correlation_*_halo, strip partitioning): input_b / backprops_b have bH >= H rows and row y of
  // input_a lines up with row y + bTop of input_b.  Plain calls: bH == H, bTop == 0.
  int bH, bTop;
  // Fused forward epilogue (fn2_correlation_forward_fused): floats between consecutive output pixels (D for a plain
  // [N,oH,oW,D] tensor, more when the volume is written into a wider concat buffer) and the LeakyReLU slope applied to
  // non-positive values (1 = none).
  int oPitch;
  float slope;
  // Fused backward (fn2_correlation_backward_fused): `gradients` is a channel window of a wider tensor with gPitch
  // floats between consecutive pixels (D for a plain [N,oH,oW,D] tensor) and, when gMask is set, every gradient is
  // multiplied by the derivative of the fused LeakyReLU, read off the sign of the stored activation at the same
  // index of gMask (same pitch): 1 where it is > 0, `slope` elsewhere (flownet_c.py:41-46 backwards).
  int gPitch;
  const float* gMask;
};

// Records a launch for fn2_launch_count() and returns the status of the launch that was just made.
int after_launch();
void note_cuda_error(cudaError_t e);

// One-time set-up that CUDA keeps PER DEVICE (cudaFuncSetAttribute: opt-in shared memory, non-portable cluster sizes).
// One flag per device ordinal; a failed attempt is not latched, and a process that moves from cuda:0 to cuda:1 runs
// the set-up again there.  Idempotent, so two threads racing on the first call of a device are harmless.
constexpr int FN2_MAX_DEVICES = 64;
struct PerDeviceOnce {
  std::atomic<int> done[FN2_MAX_DEVICES];
};
template <class F>
inline cudaError_t per_device_once(PerDeviceOnce& flags, F&& setup) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const bool tracked = dev >= 0 && dev < FN2_MAX_DEVICES;
  if (tracked && flags.done[dev].load(std::memory_order_acquire)) return cudaSuccess;
  e = setup();
  if (e == cudaSuccess && tracked) flags.done[dev].store(1, std::memory_order_release);
  return e;
}
// SM count of the current device (cached per device ordinal; 148 on B200).
int device_sm_count();
// Developer A/B switches: environment variables read ONCE, when the library is first asked (never per launch).
struct EnvSwitches {
  bool no_pair, no_split, no_persist, warp_bwd_split;
  int ds_max_ranks;   // 0 = unset
};
const EnvSwitches& env_switches();

// ---- kernels (one translation unit each) ------------------------------------------------------------------
int corr_generic_forward(const CorrGeom& g, const float* a, const float* b, float* out, cudaStream_t st);
int corr_generic_backward(const CorrGeom& g, const float* gout, const float* a, const float* b, float* ga,
                          float* gb, cudaStream_t st);

// tcgen05 kernels (tf32 / split-tf32)
bool corr_tc_supported(const CorrGeom& g, int precision);
size_t corr_tc_workspace_bytes(const CorrGeom& g, int precision, int direction);
int corr_tc_forward(const CorrGeom& g, const float* a, const float* b, float* out, int precision, void* ws,
                    size_t ws_bytes, cudaStream_t st);
int corr_tc_backward(const CorrGeom& g, const float* gout, const float* a, const float* b, float* ga, float* gb,
                     int precision, void* ws, size_t ws_bytes, cudaStream_t st);
// corr_tc_x3.cu: fp32-grade forward (split-tf32)
int corr_tc_x3_forward(const CorrGeom& g, const float* a, const float* b, float* out, cudaStream_t st);
// corr_tc_bwd.cu
bool corr_tc_bwd_supported(const CorrGeom& g, int precision);
int corr_tc_backward_launch(const CorrGeom& g, const float* gout, const float* a, const float* b, float* ga, float* gb,
                            int precision, cudaStream_t st);

int flow_warp_forward(const float* image, const float* flow, float* out, int N, int H, int W, int C,
                      cudaStream_t st);
int flow_warp_backward(const float* image, const float* flow, const float* gout, float* gimage, float* gflow, int N,
                       int H, int W, int C, cudaStream_t st);
// flow_warp_tile.cu: shared-memory window kernels for C <= 4, W % 4 == 0
bool flow_warp_tile_supported(int H, int W, int C);
int flow_warp_tile_forward(const float* image, const float* flow, float* out, int N, int H, int W, int C,
                           cudaStream_t st);
int flow_warp_tile_backward(const float* image, const float* flow, const float* gout, float* gimage, float* gflow,
                            int N, int H, int W, int C, cudaStream_t st);
int flow_warp_tile_forward_error(const float* image, const float* flow, const float* reference, float* out, float* err,
                                 int N, int H, int W, int C, cudaStream_t st);
// flow_warp.cu: warp + per-pixel L2 norm over channels of (reference - warped); any shape
int flow_warp_forward_error(const float* image, const float* flow, const float* reference, float* out, float* err, int N,
                            int H, int W, int C, cudaStream_t st);
int downsample_forward(const float* in, float* out, int N, int iH, int iW, int C, int oH, int oW, cudaStream_t st);
// downsample_multi.cu: several output sizes of one input in a single pass (host arrays of pointers / sizes), and the
// end-point-error terms of the multi-scale loss against `preds` when given (dirs, epe_sums: see fn2ops.h)
size_t multiscale_workspace_bytes(const int* out_hw, int num_scales, int N, int iH, int iW, int C);
bool multiscale_fused_supported(const int* out_hw, int num_scales, int N, int iH, int iW, int C);
int multiscale_forward(const float* in, float in_scale, float* const* outs, const float* const* preds,
                       float* const* dirs, float* epe_sums, const int* out_hw, int num_scales, int N, int iH, int iW,
                       int C, void* ws, size_t ws_bytes, cudaStream_t st, const float* const* edges = nullptr,
                       float lambda_w = 0.f);
// the plain multi-scale downsample with its scratch in the library's own device array (no workspace argument)
int multiscale_forward_noalloc(const float* in, float* const* outs, const int* out_hw, int num_scales, int N, int iH,
                               int iW, int C, cudaStream_t st);

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
// floor / ceil division for possibly negative numerators (b > 0)
__host__ __device__ inline int floor_div(int a, int b) {
  int q = a / b, r = a % b;
  return q - ((r != 0) && (r < 0));
}
__host__ __device__ inline int ceil_div_s(int a, int b) {
  int q = a / b, r = a % b;
  return q + ((r != 0) && (r > 0));
}

}  // namespace fn2
