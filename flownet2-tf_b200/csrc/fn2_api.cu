// fn2_api.cu -- the C ABI of libfn2ops.so (include/fn2ops.h): argument validation mirroring each OP_REQUIRES of the
// reference OpKernels, then dispatch to the sm_100a kernels.  No allocation, no host synchronisation.
#include <atomic>
#include <cmath>
#include <cstdlib>
#include "fn2_common.cuh"

namespace fn2 {

static std::atomic<unsigned long long> g_launches{0};
static thread_local int t_last_cuda_error = 0;

void note_cuda_error(cudaError_t e) { t_last_cuda_error = (int)e; }

int after_launch() {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    note_cuda_error(e);
    return FN2_E_CUDA;
  }
  return FN2_OK;
}

int device_sm_count() {
  static std::atomic<int> cache[FN2_MAX_DEVICES];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  const bool tracked = dev >= 0 && dev < FN2_MAX_DEVICES;
  if (tracked) {
    const int c = cache[dev].load(std::memory_order_relaxed);
    if (c > 0) return c;
  }
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms < 1) return 148;
  if (tracked) cache[dev].store(sms, std::memory_order_relaxed);
  return sms;
}

const EnvSwitches& env_switches() {
  static const EnvSwitches sw = [] {
    EnvSwitches s;
    s.no_pair = getenv("FN2_NO_PAIR") != nullptr;
    s.no_split = getenv("FN2_NO_SPLIT") != nullptr;
    s.no_persist = getenv("FN2_NO_PERSIST") != nullptr;
    s.warp_bwd_split = getenv("FN2_WARP_BWD_SPLIT") != nullptr;
    const char* e = getenv("FN2_DS_MAX_RANKS");
    s.ds_max_ranks = e ? atoi(e) : 0;
    return s;
  }();
  return sw;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// correlation_kernel.cc:23,35-57 (+ the shape fn correlation_op.cc:9-51)
static int make_geom(int N, int H, int W, int C, int k, int md, int s1, int s2, int pad, CorrGeom* g) {
  if (N < 1 || H < 1 || W < 1 || C < 1) return FN2_E_BAD_SHAPE;
  if (k < 1 || md < 0 || s1 < 1 || s2 < 1 || pad < 0) return FN2_E_BAD_PARAM;
  if (k % 2 == 0) return FN2_E_KERNEL_SIZE_EVEN;
  g->N = N; g->H = H; g->W = W; g->C = C;
  g->k = k; g->md = md; g->s1 = s1; g->s2 = s2; g->pad = pad;
  g->kr = (k - 1) / 2;
  const int border = md + g->kr;
  const int pH = H + 2 * pad, pW = W + 2 * pad;
  g->oH = (int)std::ceil((float)(pH - border * 2) / (float)s1);
  g->oW = (int)std::ceil((float)(pW - border * 2) / (float)s1);
  if (g->oH < 1) return FN2_E_OUT_HEIGHT;
  if (g->oW < 1) return FN2_E_OUT_WIDTH;
  g->r = md / s2;
  g->gw = g->r * 2 + 1;
  g->D = g->gw * g->gw;
  g->off = md - pad;
  g->bH = H;
  g->bTop = 0;
  g->oPitch = g->D;
  g->slope = 1.0f;
  g->gPitch = g->D;
  g->gMask = nullptr;
  return FN2_OK;
}

// Strip geometry: only the FlowNetC-type call (k == 1, s1 == 1, pad == md, so out rows == a
// rows) can be cut into row strips with a halo of b; a halo that is not a multiple of stride_2 would misalign the
// displacement grid.
static int add_halo(CorrGeom* g, int halo_top, int halo_bottom) {
  if (halo_top < 0 || halo_bottom < 0) return FN2_E_BAD_PARAM;
  if (g->k != 1 || g->s1 != 1 || g->pad != g->md) return FN2_E_UNSUPPORTED_GEOMETRY;
  if (halo_top % g->s2 != 0) return FN2_E_UNSUPPORTED_GEOMETRY;
  g->bTop = halo_top;
  g->bH = g->H + halo_top + halo_bottom;
  return FN2_OK;
}

}  // namespace fn2

using namespace fn2;

extern "C" {

int fn2_abi_version(void) { return FN2_ABI_VERSION; }

const char* fn2_error_string(int status) {
  switch (status) {
    case FN2_OK: return "ok";
    case FN2_E_KERNEL_SIZE_EVEN: return "kernel_size must be odd";
    case FN2_E_OUT_HEIGHT: return "Neighborhood and kernel don't fit in input height.";
    case FN2_E_OUT_WIDTH: return "Neighborhood and kernel don't fit in input width.";
    case FN2_E_BAD_SHAPE: return "tensor dimensions must be >= 1 and consistent (rank-4 NHWC)";
    case FN2_E_BAD_PARAM: return "stride_1/stride_2 must be >= 1, kernel_size >= 1, max_displacement and pad >= 0";
    case FN2_E_BAD_POINTER: return "null or misaligned device pointer (16-byte alignment required)";
    case FN2_E_UNSUPPORTED_PRECISION: return "precision not supported for these arguments";
    case FN2_E_WORKSPACE: return "workspace too small; see fn2_correlation_workspace_bytes";
    case FN2_E_DOWNSAMPLE_SIZE: return "downsample output height and width must be >= 2";
    case FN2_E_CUDA: return "CUDA error; see fn2_last_cuda_error";
    case FN2_E_NO_DEVICE: return "no sm_100 CUDA device is current";
    case FN2_E_UNSUPPORTED_GEOMETRY:
      return "unsupported geometry (row halos need kernel_size 1, stride_1 1, pad == max_displacement and "
             "halo_top % stride_2 == 0; the fused loss needs C <= 2, in_w * C % 4 == 0, in_w * C <= 1580 (a 16-row band "
             "within 100 KB of shared memory), in_h <= 16384, N <= 65535 and reductions by >= 1.5: ask "
             "fn2_multiscale_epe_supported)";
    case FN2_E_BUILTIN_SCRATCH:
      return "these shapes need more scratch than the 16 MiB array built into the library "
             "(fn2_downsample_multi_workspace_bytes > FN2_DOWNSAMPLE_MULTI_BUILTIN_SCRATCH_BYTES): use a smaller batch, "
             "or fn2_downsample_multi_forward with a caller-provided workspace";
    default: return "unknown fn2 status";
  }
}

int fn2_last_cuda_error(void) { return t_last_cuda_error; }

unsigned long long fn2_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int fn2_correlation_output_shape(int H, int W, int k, int md, int s1, int s2, int pad, int* oh, int* ow, int* oc) {
  CorrGeom g;
  const int rc = make_geom(1, H, W, 1, k, md, s1, s2, pad, &g);
  if (rc != FN2_OK) return rc;
  if (oh) *oh = g.oH;
  if (ow) *ow = g.oW;
  if (oc) *oc = g.D;
  return FN2_OK;
}

int fn2_correlation_precision_supported(int N, int H, int W, int C, int k, int md, int s1, int s2, int pad,
                                        int precision) {
  CorrGeom g;
  const int rc = make_geom(N, H, W, C, k, md, s1, s2, pad, &g);
  if (rc != FN2_OK) return rc;
  if (precision == FN2_PREC_FP32) return 1;
  if (precision != FN2_PREC_TF32 && precision != FN2_PREC_TF32X3) return 0;
  return corr_tc_supported(g, precision) ? 1 : 0;
}

size_t fn2_correlation_workspace_bytes(int N, int H, int W, int C, int k, int md, int s1, int s2, int pad,
                                       int precision, int direction) {
  CorrGeom g;
  if (make_geom(N, H, W, C, k, md, s1, s2, pad, &g) != FN2_OK) return 0;
  if (precision == FN2_PREC_FP32) return 0;
  if (!corr_tc_supported(g, precision)) return 0;
  return corr_tc_workspace_bytes(g, precision, direction);
}

static int correlation_forward(const CorrGeom& g, const float* a, const float* b, float* out, int precision, void* ws,
                               size_t ws_bytes, void* stream);
static int correlation_backward(const CorrGeom& g, const float* gout, const float* a, const float* b, float* ga,
                                float* gb, int precision, void* ws, size_t ws_bytes, void* stream);

int fn2_correlation_forward(const float* a, const float* b, float* out, int N, int H, int W, int C, int k, int md,
                            int s1, int s2, int pad, int precision, void* ws, size_t ws_bytes, void* stream) {
  CorrGeom g;
  const int rc = make_geom(N, H, W, C, k, md, s1, s2, pad, &g);
  if (rc != FN2_OK) return rc;
  return correlation_forward(g, a, b, out, precision, ws, ws_bytes, stream);
}

int fn2_correlation_forward_fused(const float* a, const float* b, float* out, int N, int H, int W, int C, int k, int md,
                                  int s1, int s2, int pad, int out_pixel_pitch, float leaky_slope, int precision,
                                  void* ws, size_t ws_bytes, void* stream) {
  CorrGeom g;
  const int rc = make_geom(N, H, W, C, k, md, s1, s2, pad, &g);
  if (rc != FN2_OK) return rc;
  if (out_pixel_pitch < g.D) return FN2_E_BAD_SHAPE;
  // The backward of the fused activation recovers its mask from the sign of the stored output, which needs slope > 0
  // (the reference's LeakyReLU is 0.1, utils.py:401-405); NaN is rejected by the same test.
  if (!(leaky_slope > 0.f)) return FN2_E_BAD_PARAM;
  g.oPitch = out_pixel_pitch;
  g.slope = leaky_slope;
  return correlation_forward(g, a, b, out, precision, ws, ws_bytes, stream);
}

int fn2_correlation_backward_fused(const float* gout, const float* activations, const float* a, const float* b, float* ga,
                                   float* gb, int N, int H, int W, int C, int k, int md, int s1, int s2, int pad,
                                   int pixel_pitch, float leaky_slope, int precision, void* ws, size_t ws_bytes,
                                   void* stream) {
  CorrGeom g;
  const int rc = make_geom(N, H, W, C, k, md, s1, s2, pad, &g);
  if (rc != FN2_OK) return rc;
  if (pixel_pitch < g.D) return FN2_E_BAD_SHAPE;
  if (!(leaky_slope > 0.f)) return FN2_E_BAD_PARAM;   // the mask is the sign of the stored activation
  if (!activations) return FN2_E_BAD_POINTER;
  g.gPitch = pixel_pitch;
  g.gMask = activations;
  g.slope = leaky_slope;
  // a channel window of a wider tensor starts at an arbitrary channel: 4-byte alignment is all these two need
  return correlation_backward(g, gout, a, b, ga, gb, precision, ws, ws_bytes, stream);
}

int fn2_correlation_forward_halo(const float* a, const float* b, float* out, int N, int H, int W, int C, int k, int md,
                                 int s1, int s2, int pad, int halo_top, int halo_bottom, int precision, void* ws,
                                 size_t ws_bytes, void* stream) {
  CorrGeom g;
  int rc = make_geom(N, H, W, C, k, md, s1, s2, pad, &g);
  if (rc == FN2_OK) rc = add_halo(&g, halo_top, halo_bottom);
  if (rc != FN2_OK) return rc;
  return correlation_forward(g, a, b, out, precision, ws, ws_bytes, stream);
}

int fn2_correlation_backward_halo(const float* gout, const float* a, const float* b, float* ga, float* gb, int N,
                                  int H, int W, int C, int k, int md, int s1, int s2, int pad, int halo_top,
                                  int halo_bottom, int precision, void* ws, size_t ws_bytes, void* stream) {
  CorrGeom g;
  int rc = make_geom(N, H, W, C, k, md, s1, s2, pad, &g);
  if (rc == FN2_OK) rc = add_halo(&g, halo_top, halo_bottom);
  if (rc != FN2_OK) return rc;
  return correlation_backward(g, gout, a, b, ga, gb, precision, ws, ws_bytes, stream);
}

static int correlation_forward(const CorrGeom& g, const float* a, const float* b, float* out, int precision, void* ws,
                               size_t ws_bytes, void* stream) {
  if (!a || !b || !out || !aligned16(a) || !aligned16(b) || !aligned16(out)) return FN2_E_BAD_POINTER;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (precision) {
    case FN2_PREC_FP32:
      return corr_generic_forward(g, a, b, out, st);   // handles row halos too
    case FN2_PREC_TF32:
    case FN2_PREC_TF32X3:
      if (!corr_tc_supported(g, precision)) return FN2_E_UNSUPPORTED_PRECISION;
      if (ws_bytes < corr_tc_workspace_bytes(g, precision, 0) || (ws_bytes && !ws)) return FN2_E_WORKSPACE;
      return corr_tc_forward(g, a, b, out, precision, ws, ws_bytes, st);
    default:
      return FN2_E_UNSUPPORTED_PRECISION;
  }
}

int fn2_correlation_backward(const float* gout, const float* a, const float* b, float* ga, float* gb, int N, int H,
                             int W, int C, int k, int md, int s1, int s2, int pad, int precision, void* ws,
                             size_t ws_bytes, void* stream) {
  CorrGeom g;
  const int rc = make_geom(N, H, W, C, k, md, s1, s2, pad, &g);
  if (rc != FN2_OK) return rc;
  return correlation_backward(g, gout, a, b, ga, gb, precision, ws, ws_bytes, stream);
}

static int correlation_backward(const CorrGeom& g, const float* gout, const float* a, const float* b, float* ga,
                                float* gb, int precision, void* ws, size_t ws_bytes, void* stream) {
  if (!gout || !a || !b || !ga || !gb) return FN2_E_BAD_POINTER;
  if (!aligned16(a) || !aligned16(b) || !aligned16(ga) || !aligned16(gb)) return FN2_E_BAD_POINTER;
  if (g.gMask ? ((reinterpret_cast<uintptr_t>(gout) | reinterpret_cast<uintptr_t>(g.gMask)) & 3u) != 0 : !aligned16(gout))
    return FN2_E_BAD_POINTER;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (precision) {
    case FN2_PREC_FP32:
      return corr_generic_backward(g, gout, a, b, ga, gb, st);
    case FN2_PREC_TF32:
    case FN2_PREC_TF32X3:
      if (!corr_tc_supported(g, precision)) return FN2_E_UNSUPPORTED_PRECISION;
      if (ws_bytes < corr_tc_workspace_bytes(g, precision, 1) || (ws_bytes && !ws)) return FN2_E_WORKSPACE;
      return corr_tc_backward(g, gout, a, b, ga, gb, precision, ws, ws_bytes, st);
    default:
      return FN2_E_UNSUPPORTED_PRECISION;
  }
}

int fn2_flow_warp_forward(const float* image, const float* flow, float* out, int N, int H, int W, int C,
                          void* stream) {
  if (N < 1 || H < 1 || W < 1 || C < 1) return FN2_E_BAD_SHAPE;
  if (!image || !flow || !out || !aligned16(image) || !aligned16(flow) || !aligned16(out)) return FN2_E_BAD_POINTER;
  return flow_warp_forward(image, flow, out, N, H, W, C, static_cast<cudaStream_t>(stream));
}

int fn2_flow_warp_error_forward(const float* image, const float* flow, const float* reference, float* out, float* err,
                                int N, int H, int W, int C, void* stream) {
  if (N < 1 || H < 1 || W < 1 || C < 1) return FN2_E_BAD_SHAPE;
  if (!image || !flow || !reference || !out || !err) return FN2_E_BAD_POINTER;
  if (!aligned16(image) || !aligned16(flow) || !aligned16(reference) || !aligned16(out) || !aligned16(err))
    return FN2_E_BAD_POINTER;
  return flow_warp_forward_error(image, flow, reference, out, err, N, H, W, C, static_cast<cudaStream_t>(stream));
}

int fn2_flow_warp_backward(const float* image, const float* flow, const float* gout, float* gimage, float* gflow,
                           int N, int H, int W, int C, void* stream) {
  if (N < 1 || H < 1 || W < 1 || C < 1) return FN2_E_BAD_SHAPE;
  if (!image || !flow || !gout || !gimage || !gflow) return FN2_E_BAD_POINTER;
  if (!aligned16(image) || !aligned16(flow) || !aligned16(gout) || !aligned16(gimage) || !aligned16(gflow))
    return FN2_E_BAD_POINTER;
  return flow_warp_backward(image, flow, gout, gimage, gflow, N, H, W, C, static_cast<cudaStream_t>(stream));
}

int fn2_downsample_forward(const float* in, float* out, int N, int iH, int iW, int C, int oH, int oW, void* stream) {
  if (N < 1 || iH < 1 || iW < 1 || C < 1 || oH < 1 || oW < 1) return FN2_E_BAD_SHAPE;
  if (oH < 2 || oW < 2) return FN2_E_DOWNSAMPLE_SIZE;
  if (!in || !out || !aligned16(in) || !aligned16(out)) return FN2_E_BAD_POINTER;
  return downsample_forward(in, out, N, iH, iW, C, oH, oW, static_cast<cudaStream_t>(stream));
}

static int check_scales(const int* out_hw, int num_scales, int N, int iH, int iW, int C) {
  if (N < 1 || iH < 1 || iW < 1 || C < 1 || num_scales < 1) return FN2_E_BAD_SHAPE;
  if (!out_hw) return FN2_E_BAD_POINTER;
  for (int s = 0; s < num_scales; ++s) {
    if (out_hw[2 * s] < 1 || out_hw[2 * s + 1] < 1) return FN2_E_BAD_SHAPE;
    if (out_hw[2 * s] < 2 || out_hw[2 * s + 1] < 2) return FN2_E_DOWNSAMPLE_SIZE;
  }
  return FN2_OK;
}

size_t fn2_downsample_multi_workspace_bytes(const int* out_hw, int num_scales, int N, int iH, int iW, int C) {
  if (check_scales(out_hw, num_scales, N, iH, iW, C) != FN2_OK) return 0;
  return multiscale_workspace_bytes(out_hw, num_scales, N, iH, iW, C);
}

int fn2_downsample_multi_forward(const float* in, float* const* outs, const int* out_hw, int num_scales, int N, int iH,
                                 int iW, int C, void* ws, size_t ws_bytes, void* stream) {
  const int rc = check_scales(out_hw, num_scales, N, iH, iW, C);
  if (rc != FN2_OK) return rc;
  if (!in || !outs || !aligned16(in)) return FN2_E_BAD_POINTER;
  for (int s = 0; s < num_scales; ++s)
    if (!outs[s] || !aligned16(outs[s])) return FN2_E_BAD_POINTER;
  if (ws_bytes && (!ws || !aligned16(ws))) return FN2_E_BAD_POINTER;
  return multiscale_forward(in, 1.0f, outs, nullptr, nullptr, nullptr, out_hw, num_scales, N, iH, iW, C, ws, ws_bytes,
                            static_cast<cudaStream_t>(stream));
}

// No workspace argument and no allocation: the scratch is the library's own device array (downsample_multi.cu), for
// callers that have no device memory to give.  The first call on a device also finds out where that array lives there.
int fn2_downsample_multi_forward_noalloc(const float* in, float* const* outs, const int* out_hw, int num_scales, int N,
                                         int iH, int iW, int C, void* stream) {
  const int rc = check_scales(out_hw, num_scales, N, iH, iW, C);
  if (rc != FN2_OK) return rc;
  if (!in || !outs || !aligned16(in)) return FN2_E_BAD_POINTER;
  for (int s = 0; s < num_scales; ++s)
    if (!outs[s] || !aligned16(outs[s])) return FN2_E_BAD_POINTER;
  return multiscale_forward_noalloc(in, outs, out_hw, num_scales, N, iH, iW, C, static_cast<cudaStream_t>(stream));
}

int fn2_multiscale_epe_supported(const int* out_hw, int num_scales, int N, int iH, int iW, int C) {
  const int rc = check_scales(out_hw, num_scales, N, iH, iW, C);
  if (rc != FN2_OK) return rc;
  return multiscale_fused_supported(out_hw, num_scales, N, iH, iW, C) ? 1 : 0;
}

int fn2_multiscale_epe_forward(const float* flow, float flow_scale, const float* const* preds, const int* out_hw,
                               int num_scales, int N, int iH, int iW, int C, float* const* downsampled,
                               float* const* residual_dir, float* epe_sums, void* ws, size_t ws_bytes, void* stream) {
  const int rc = check_scales(out_hw, num_scales, N, iH, iW, C);
  if (rc != FN2_OK) return rc;
  if (!(flow_scale == flow_scale)) return FN2_E_BAD_PARAM;
  if (!flow || !preds || !epe_sums || !aligned16(flow)) return FN2_E_BAD_POINTER;
  for (int s = 0; s < num_scales; ++s) {
    if (!preds[s] || !aligned16(preds[s])) return FN2_E_BAD_POINTER;
    if (downsampled && downsampled[s] && !aligned16(downsampled[s])) return FN2_E_BAD_POINTER;
    if (residual_dir && residual_dir[s] && !aligned16(residual_dir[s])) return FN2_E_BAD_POINTER;
  }
  if (ws_bytes && (!ws || !aligned16(ws))) return FN2_E_BAD_POINTER;
  return multiscale_forward(flow, flow_scale, downsampled, preds, residual_dir, epe_sums, out_hw, num_scales, N, iH, iW,
                            C, ws, ws_bytes, static_cast<cudaStream_t>(stream));
}

int fn2_multiscale_epe_edges_forward(const float* flow, float flow_scale, const float* const* preds,
                                     const float* const* edges, float lambda_w, const int* out_hw, int num_scales,
                                     int N, int iH, int iW, int C, float* const* downsampled,
                                     float* const* residual_dir, float* epe_sums, void* ws, size_t ws_bytes,
                                     void* stream) {
  const int rc = check_scales(out_hw, num_scales, N, iH, iW, C);
  if (rc != FN2_OK) return rc;
  if (!(flow_scale == flow_scale) || !(lambda_w == lambda_w)) return FN2_E_BAD_PARAM;
  if (!flow || !preds || !edges || !epe_sums || !aligned16(flow)) return FN2_E_BAD_POINTER;
  for (int s = 0; s < num_scales; ++s) {
    if (!preds[s] || !aligned16(preds[s]) || !edges[s]) return FN2_E_BAD_POINTER;
    if (downsampled && downsampled[s] && !aligned16(downsampled[s])) return FN2_E_BAD_POINTER;
    if (residual_dir && residual_dir[s] && !aligned16(residual_dir[s])) return FN2_E_BAD_POINTER;
  }
  if (ws_bytes && (!ws || !aligned16(ws))) return FN2_E_BAD_POINTER;
  return multiscale_forward(flow, flow_scale, downsampled, preds, residual_dir, epe_sums, out_hw, num_scales, N, iH, iW,
                            C, ws, ws_bytes, static_cast<cudaStream_t>(stream), edges, lambda_w);
}

}  // extern "C"
