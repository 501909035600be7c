/*
 * fn2ops.h -- C ABI of libfn2ops.so: the B200 (sm_100a) replacement for the three TensorFlow custom-op
 * libraries of fperezgamonal/flownet2-tf (correlation.so, flow_warp.so, downsample.so).
 *
 * Every entry point names the reference interface it replaces (paths relative to the reference repo).  The
 * reference binds its ops with tf.load_op_library + OpKernel::Compute; this library is bound with ctypes (see
 * INTEGRATION.md for the stub a maintainer adds to src/correlation.py, src/flow_warp.py, src/downsample.py).
 *
 * Conventions
 *   - All tensors are NHWC, contiguous, float32, resident in device memory of the current CUDA device.
 *   - The caller owns every buffer (inputs, outputs, workspace).  The library allocates nothing, keeps no
 *     pointer after returning, and fully defines every output element (no pre-zeroing needed).
 *   - Every call only enqueues work on `stream` (a cudaStream_t passed as void*; NULL = legacy default
 *     stream) and returns without synchronising; calls are re-entrant and CUDA-graph capturable.
 *     (fn2_downsample_multi_forward_noalloc keeps its scratch inside the library and states its own rules.)
 *   - Return value: FN2_OK (0) or a negative fn2_status.  Nothing throws, prints or exits.
 *   - Element counts above 2^31 are supported (64-bit indexing); each single dimension must fit in int.
 */
#ifndef FN2OPS_H_
#define FN2OPS_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 2: fn2_downsample_multi_forward takes (workspace, workspace_bytes) before stream, precision value 2 (bf16) is
 *    rejected; fn2_downsample_multi_forward_noalloc, fn2_multiscale_epe_supported and
 *    fn2_multiscale_epe_edges_forward added.  A caller built against
 *    version 1 must not bind this library: check fn2_abi_version() == FN2_ABI_VERSION after loading. */
#define FN2_ABI_VERSION 2

#if defined(__GNUC__)
#define FN2_API __attribute__((visibility("default")))
#else
#define FN2_API
#endif

typedef enum fn2_status {
  FN2_OK = 0,
  FN2_E_KERNEL_SIZE_EVEN = -1, /* "kernel_size must be odd"            correlation_kernel.cc:23            */
  FN2_E_OUT_HEIGHT = -2,       /* "Neighborhood and kernel don't fit in input height."  correlation_kernel.cc:50 */
  FN2_E_OUT_WIDTH = -3,        /* "Neighborhood and kernel don't fit in input width."   correlation_kernel.cc:52 */
  FN2_E_BAD_SHAPE = -4,        /* a dimension < 1, or shapes that the op's shape fn would reject         */
  FN2_E_BAD_PARAM = -5,        /* stride < 1, max_displacement < 0, pad < 0                              */
  FN2_E_BAD_POINTER = -6,      /* NULL or misaligned (float32 tensors must be 16-byte aligned)           */
  FN2_E_UNSUPPORTED_PRECISION = -7,
  FN2_E_WORKSPACE = -8,        /* workspace smaller than fn2_correlation_workspace_bytes()              */
  FN2_E_DOWNSAMPLE_SIZE = -9,  /* output height or width < 2: the reference divides by zero there
                                  (downsample_kernel_gpu.cu.cc:40-41,92-93); rejected instead           */
  FN2_E_CUDA = -10,            /* a CUDA runtime/driver call failed; see fn2_last_cuda_error()          */
  FN2_E_NO_DEVICE = -11,       /* no sm_100 device is current                                           */
  FN2_E_UNSUPPORTED_GEOMETRY = -12,/* *_halo entry points: need kernel_size 1, stride_1 1, pad == max_displacement,
                                  halo_top a multiple of stride_2; fn2_multiscale_epe_forward: shapes the
                                  fused multi-scale kernels do not take (fn2_multiscale_epe_supported)  */
  FN2_E_BUILTIN_SCRATCH = -13  /* fn2_downsample_multi_forward_noalloc: the shapes need more scratch than the
                                  library's built-in array holds (FN2_DOWNSAMPLE_MULTI_BUILTIN_SCRATCH_BYTES)  */
} fn2_status;

/* Arithmetic used by the correlation contraction.  Inputs and outputs are float32 in every mode.
 * The reference op is float32 only (correlation_op.cc:54-61); FP32 and TF32X3 meet its tolerance (1e-5 relative +
 * 1e-6 absolute against a float64 evaluation), TF32 trades accuracy (2.5e-3 of the tensor's maximum) for speed.
 * (TF32X3, gradients only: the tensor core truncates when adding into its accumulator and wide images take more
 * accumulation steps; at W = 480 with C = 32 and unit-variance data about 1 element in 10^4 of backprops_a reaches
 * 4.4e-6 absolute.  FlowNetC's C = 256 stays inside the tolerance at every width tested, 4K included.)
 * The tensor-core modes cover the FlowNetC call shape (kernel_size 1, stride_1 1, pad == max_displacement,
 * max_displacement / stride_2 == 10, stride_2 in {1, 2}, C a multiple of 32 and <= 256: flownet_c.py:40); ask
 * fn2_correlation_precision_supported().  Value 2 is reserved (a bf16 mode is not built). */
typedef enum fn2_precision {
  FN2_PREC_FP32 = 0,   /* fp32 FMA on the CUDA cores, any geometry                                          */
  FN2_PREC_TF32 = 1,   /* tcgen05 kind::tf32, fp32 accumulate in TMEM                                      */
  FN2_PREC_TF32X3 = 3  /* tcgen05 split-tf32: operands split hi + lo, the partial products (hi*hi, hi*lo,
                          lo*hi; the forward also lo*lo) accumulated in fp32: fp32-grade on tensor cores   */
} fn2_precision;

FN2_API int fn2_abi_version(void);
FN2_API const char* fn2_error_string(int status);
/* cudaError_t of the last failing CUDA call made by this thread through the library (0 if none). */
FN2_API int fn2_last_cuda_error(void);

/* ---------------------------------------------------------------------------------------------------------
 * Correlation.  Replaces REGISTER_OP("Correlation") + CorrelationKernel<GPUDevice>::Compute
 * (src/ops/correlation/correlation_op.cc:9-62, correlation_kernel.cc:13-136) and the Pad/CorrelateData
 * launchers beneath it (pad.cu.cc:46-74, correlation_kernel.cu.cc:121-150).  No padded temporaries are made.
 *
 *   out[n,yo,xo,(p+r)*gw+(o+r)] = 1/(k*k*C) * sum_{j,i<k} sum_c  A[n, yo*s1+md+j,      xo*s1+md+i,      c]
 *                                                              * B[n, yo*s1+md+j+s2*p, xo*s1+md+i+s2*o, c]
 *   with A,B = a,b zero-padded by `pad`, r = md / s2 (integer), gw = 2r+1, p,o in [-r, r].
 * --------------------------------------------------------------------------------------------------------- */

/* Shape function (correlation_op.cc:9-51, correlation_kernel.cc:35-57): out = [N, *out_h, *out_w, *out_c]. */
FN2_API int fn2_correlation_output_shape(int H, int W, int kernel_size, int max_displacement, int stride_1, int stride_2,
                                 int pad, int* out_h, int* out_w, int* out_c);

/* 1 if `precision` can run this call, 0 if not (fn2_correlation_* would return FN2_E_UNSUPPORTED_PRECISION), or a
 * negative fn2_status for arguments the op itself rejects.  FN2_PREC_FP32 is supported for every valid call. */
FN2_API int fn2_correlation_precision_supported(int N, int H, int W, int C, int kernel_size, int max_displacement,
                                        int stride_1, int stride_2, int pad, int precision);

/* Scratch bytes needed by the chosen precision (0 for FN2_PREC_FP32).  direction: 0 = forward, 1 = backward. */
FN2_API size_t fn2_correlation_workspace_bytes(int N, int H, int W, int C, int kernel_size, int max_displacement,
                                       int stride_1, int stride_2, int pad, int precision, int direction);

FN2_API int fn2_correlation_forward(const float* input_a, const float* input_b, float* output, int N, int H, int W, int C,
                            int kernel_size, int max_displacement, int stride_1, int stride_2, int pad,
                            int precision, void* workspace, size_t workspace_bytes, void* stream);

/* Correlation with the two steps that follow it in FlowNetC fused into the store:
 *   cc_relu = LeakyReLU(correlation(conv_a_3, conv_b_3, ...))            flownet_c.py:40-41, utils.py:401-405
 *   net     = tf.concat([netA_conv, cc_relu], axis=3)                    flownet_c.py:44-46
 * `output` points at the first correlation channel of pixel (0,0,0) INSIDE the wider buffer (e.g. concat + 32 for the
 * 32 conv_redir channels), `out_pixel_pitch` is that buffer's channel count (473), and values <= 0 are multiplied by
 * `leaky_slope` (0.1; 1 = no activation).  Only the D correlation channels of every pixel are written.  With
 * out_pixel_pitch == D and leaky_slope == 1 this is fn2_correlation_forward. */
FN2_API int fn2_correlation_forward_fused(const float* input_a, const float* input_b, float* output, int N, int H, int W,
                                  int C, int kernel_size, int max_displacement, int stride_1, int stride_2, int pad,
                                  int out_pixel_pitch, float leaky_slope, int precision, void* workspace,
                                  size_t workspace_bytes, void* stream);

/* The backward of fn2_correlation_forward_fused, with the two steps that precede CorrelationGrad in the reference's
 * gradient graph fused into the kernel's gradient gather:
 *   d cc_relu = slice(d net, channels [offset, offset + D))                  gradient of tf.concat, flownet_c.py:44-46
 *   d cc      = d cc_relu * (cc_relu > 0 ? 1 : leaky_slope)                  gradient of LeakyReLU,  flownet_c.py:41
 * `gradients` points at the first correlation channel of pixel (0,0,0) inside the gradient of the concat buffer,
 * `activations` at the same place inside the buffer the forward wrote (the mask is the sign of the stored value, which
 * needs leaky_slope > 0); both have `pixel_pitch` floats per pixel and only need 4-byte alignment.  No masked
 * 441-channel copy is materialised.  With pixel_pitch == D and leaky_slope == 1 this is fn2_correlation_backward. */
FN2_API int fn2_correlation_backward_fused(const float* gradients, const float* activations, const float* input_a,
                                   const float* input_b, float* backprops_a, float* backprops_b, int N, int H, int W,
                                   int C, int kernel_size, int max_displacement, int stride_1, int stride_2, int pad,
                                   int pixel_pitch, float leaky_slope, int precision, void* workspace,
                                   size_t workspace_bytes, void* stream);

/* Replaces REGISTER_OP("CorrelationGrad") + CorrelationGradKernel<GPUDevice>::Compute
 * (correlation_op.cc:64-82, correlation_grad_kernel.cc:15-159) and CorrelationGradA/B
 * (correlation_grad_kernel.cu.cc:191-259).  One launch set for the whole batch; no atomics; both gradients
 * are fully written. */
FN2_API int fn2_correlation_backward(const float* gradients, const float* input_a, const float* input_b, float* backprops_a,
                             float* backprops_b, int N, int H, int W, int C, int kernel_size, int max_displacement,
                             int stride_1, int stride_2, int pad, int precision, void* workspace,
                             size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Row-strip variants (4K inputs cut into horizontal strips, one per GPU).  The reference has
 * no such entry point: its CorrelationKernel::Compute (correlation_kernel.cc:13-136) always sees whole images; these
 * compute exactly the rows of that result which belong to one strip.
 *
 *   input_a, output, gradients, backprops_a : the strip's own H rows
 *   input_b, backprops_b                    : H + halo_top + halo_bottom rows; row y of input_a lines up with row
 *                                             y + halo_top of input_b.  Rows beyond either end of input_b are the op's
 *                                             zero padding (a strip at the image border passes halo 0 on that side).
 * backprops_b is written for ALL rows of the tall tensor: its halo rows are the partial sums that belong to the
 * neighbouring strips (the caller sends them back and adds them).  Requires kernel_size 1, stride_1 1,
 * pad == max_displacement (FlowNetC's call, flownet_c.py:40) and halo_top % stride_2 == 0.
 * --------------------------------------------------------------------------------------------------------- */
FN2_API int fn2_correlation_forward_halo(const float* input_a, const float* input_b, float* output, int N, int H, int W,
                                 int C, int kernel_size, int max_displacement, int stride_1, int stride_2, int pad,
                                 int halo_top, int halo_bottom, int precision, void* workspace,
                                 size_t workspace_bytes, void* stream);

FN2_API int fn2_correlation_backward_halo(const float* gradients, const float* input_a, const float* input_b,
                                  float* backprops_a, float* backprops_b, int N, int H, int W, int C,
                                  int kernel_size, int max_displacement, int stride_1, int stride_2, int pad,
                                  int halo_top, int halo_bottom, int precision, void* workspace,
                                  size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * FlowWarp.  Replaces REGISTER_OP("FlowWarp") + FlowWarpKernel<GPUDevice>::Compute
 * (src/ops/flow_warp/flow_warp_op.cc:6-11, flow_warp.cc:13-47, launcher flow_warp.cu.cc:99-127).
 * image [N,H,W,C], flow [N,H,W,2] (channel 0 = x, 1 = y), output [N,H,W,C].
 * --------------------------------------------------------------------------------------------------------- */
FN2_API int fn2_flow_warp_forward(const float* image, const float* flow, float* output, int N, int H, int W, int C,
                          void* stream);

/* FlowWarp with the brightness error that every stacked net computes right after it, in the same pass (SURVEY.md
 * section 8 f-2):  warped = flow_warp(input_b, flow);  error = sqrt(reduce_sum(square(input_a - warped), axis=3))
 * (flownet_cs.py:21-27, flownet_css.py:21-27, flownet2.py:33-39).  reference = input_a [N,H,W,C];
 * output [N,H,W,C] = warped, error [N,H,W,1].  Forward only: in the reference these nodes sit behind frozen nets and
 * data inputs and are never differentiated. */
FN2_API int fn2_flow_warp_error_forward(const float* image, const float* flow, const float* reference, float* output,
                                float* error, int N, int H, int W, int C, void* stream);

/* Replaces REGISTER_OP("FlowWarpGrad") + FlowWarpGradKernel<GPUDevice>::Compute
 * (flow_warp_op.cc:13-22, flow_warp_grad.cc:13-56, launcher flow_warp_grad.cu.cc:90-123) with the two outputs the
 * op declares (the reference allocates both at index 0, flow_warp_grad.cc:36-37, and so cannot run).
 * image_grad is zero-filled inside, on `stream`.  image_grad accumulates with fp32 atomics: run-to-run results
 * may differ in the last bits, exactly as in the reference. */
FN2_API int fn2_flow_warp_backward(const float* image, const float* flow, const float* gradients, float* image_grad,
                           float* flow_grad, int N, int H, int W, int C, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Downsample.  Replaces REGISTER_OP("Downsample") + DownsampleKernel<GPUDevice>::Compute
 * (src/ops/downsample/downsample_op.cc:11-28, downsample_kernel.cc:13-46, launcher
 * downsample_kernel_gpu.cu.cc:80-104).  input [N,in_h,in_w,C] -> output [N,out_h,out_w,C].
 * --------------------------------------------------------------------------------------------------------- */
FN2_API int fn2_downsample_forward(const float* input, float* output, int N, int in_h, int in_w, int C, int out_h,
                           int out_w, void* stream);

/* Multi-scale downsample.  Replaces the five consecutive downsample() calls that every loss of
 * the reference makes on the same ground-truth flow (flownet_c.py:135-160, flownet_s.py:128-153, flownet_sd.py:129-154;
 * edge maps likewise in the HFEM losses, utils.py:227-339): outputs[s] = downsample(input, [out_hw[2s], out_hw[2s+1]])
 * for s < num_scales, every scale computed directly from `input` with the arithmetic of fn2_downsample_forward, the
 * input read from HBM once.  `outputs` (num_scales device pointers) and `out_hw` (2*num_scales ints) are HOST arrays,
 * consumed before the call returns.  `workspace`: fn2_downsample_multi_workspace_bytes() bytes of device memory. */
FN2_API int fn2_downsample_multi_forward(const float* input, float* const* outputs, const int* out_hw, int num_scales,
                                 int N, int in_h, int in_w, int C, void* workspace, size_t workspace_bytes,
                                 void* stream);

/* Scratch bytes of the two multi-scale entry points (tables that depend on the shapes only, and the per-tile partial
 * sums of the band pass).  0 when every scale goes through the single-scale kernels (C > 2, upsampling, odd widths). */
FN2_API size_t fn2_downsample_multi_workspace_bytes(const int* out_hw, int num_scales, int N, int in_h, int in_w, int C);

/* fn2_downsample_multi_forward for callers that have no device buffer to hand over as `workspace`.  It has no
 * counterpart in the reference (whose five downsample() calls need no scratch at all); it exists for the binding
 * INTEGRATION.md section 1 describes.  A ctypes stub living inside a TensorFlow process sees only the tensors the
 * graph gives it: TensorFlow's allocator has already reserved (by default nearly all of) the GPU's memory for itself,
 * so a cudaMalloc from the stub fails or eats into what is left for the driver, and allocate_temp is out of reach from
 * Python.  The same holds for any host that gives a plug-in its inputs and outputs and nothing else.  Such callers
 * cannot use fn2_downsample_multi_forward, which asks them for fn2_downsample_multi_workspace_bytes() of scratch.
 *
 * This variant takes the same inputs and outputs and runs the same kernels, with the scratch (weight tables, per-band
 * partial sums, a small control block) placed in an array that is part of the library image itself:
 * FN2_DOWNSAMPLE_MULTI_BUILTIN_SCRATCH_BYTES of device memory on every GPU that loads libfn2ops.so, held until the
 * process exits.  It calls no cudaMalloc / cudaFree and nothing that allocates on its behalf.
 *
 *   - Limit: calls whose fn2_downsample_multi_workspace_bytes() exceeds the array are refused with
 *     FN2_E_BUILTIN_SCRATCH before anything is enqueued (16 MiB covers the loss sizes of flownet_c.py:135-160 on
 *     384 x 512 x 2 flows up to a batch of 18; shapes that go through the single-scale kernels need no scratch
 *     and always fit).
 *   - First call on each device: the library looks up where the array lives on the CURRENT device (a driver query that
 *     may also load the kernels' module) and remembers it per device ordinal.  Make that first call during set-up, on
 *     the device that will be used, before any stream capture is begun.
 *   - One array per device, shared by every call: two calls that can overlap on the GPU (different streams, or different
 *     host threads on different streams) race on it and corrupt each other's results.  Issue them on one stream, or
 *     order them with events.  Calls on one stream need nothing: the array keeps whatever the previous call left, and
 *     every call rewrites the tables and partial sums it reads and resets its control word, in stream order. */
#define FN2_DOWNSAMPLE_MULTI_BUILTIN_SCRATCH_BYTES (16u * 1024u * 1024u)
FN2_API int fn2_downsample_multi_forward_noalloc(const float* input, float* const* outputs, const int* out_hw,
                                         int num_scales, int N, int in_h, int in_w, int C, void* stream);

/* The data term of the multi-scale training loss in the same pass (second half).  Replaces,
 * for every scale s, the graph nodes of flownet_c.py:127-160 + utils.py:209-224:
 *     flow_s   = downsample(flow * flow_scale, [out_hw[2s], out_hw[2s+1]])          (flow * 0.05: flownet_c.py:128)
 *     epe_s    = sqrt(reduce_sum(square(predictions[s] - flow_s), axis=3))
 *     epe_sums[s] = reduce_sum(epe_s) / N                                             (average_endpoint_error)
 * flow [N,in_h,in_w,C] is multiplied by flow_scale on load (a rounded fp32 product, exactly the materialised tensor of
 * the reference); predictions[s] is [N,out_h,out_w,C].  Optional outputs (the array, or single entries, may be NULL):
 * downsampled[s] = flow_s, residual_dir[s] = (predictions[s] - flow_s) / epe_s, the factor the gradient of epe_sums[s]
 * with respect to predictions[s] is made of (0/0 = NaN where the error is exactly 0, as the gradient of sqrt there).
 * epe_sums is a DEVICE array of num_scales floats; the sums are formed in a fixed order (bit-reproducible).
 * `predictions`, `downsampled`, `residual_dir` and `out_hw` are HOST arrays consumed before the call returns.
 * Needs the fused kernels for every scale: FN2_E_UNSUPPORTED_GEOMETRY otherwise (there is no slower path inside this
 * entry point; ask fn2_multiscale_epe_supported first and build the terms from fn2_downsample_multi_forward when it
 * says 0, as flownet2_tf_b200.downsample.multiscale_epe does).  The fused kernels take
 *     C <= 2;  in_w * C a multiple of 4 and <= 1580 (a band of 16 input rows must fit 100 KB of shared memory: flows
 *     up to 790 px wide at C = 2 -- FlyingChairs' 512 yes, Sintel's 1024 and KITTI's 1242 no);  in_h <= 16384;
 *     N <= 65535;  every scale a reduction by >= 1.5 in both directions. */
FN2_API int fn2_multiscale_epe_forward(const float* flow, float flow_scale, const float* const* predictions,
                               const int* out_hw, int num_scales, int N, int in_h, int in_w, int C,
                               float* const* downsampled, float* const* residual_dir, float* epe_sums,
                               void* workspace, size_t workspace_bytes, void* stream);

/* fn2_multiscale_epe_forward with the edge weighting of the reference's HFEM loss in mode 'edges'
 * (average_endpoint_error_hfem, utils.py:227-339, branch :326-334; callers flownet_s_interp.py:170-239, which
 * downsample the edge map to every scale first, :177,191,205,219,233):
 *     epe_sums[s] = reduce_sum(epe_s + lambda_w * (epe_s * edges[s])) / N
 * edges[s] is a device tensor [N,out_h,out_w,1] -- the edge map ALREADY downsampled to scale s (one
 * fn2_downsample_multi_forward call on the [N,in_h,in_w,1] map gives all of them); `edges` itself is a HOST array like
 * `predictions`.  residual_dir[s] then holds (1 + lambda_w * edges[s]) * (predictions[s] - flow_s) / epe_s, again the
 * whole factor of the gradient with respect to predictions[s].  Everything else as fn2_multiscale_epe_forward (same
 * shape limits, same workspace).  The 'hard' mode (top-k of the error map, :268-324) is not a fused kernel: the host
 * side selects on the per-pixel errors. */
FN2_API int fn2_multiscale_epe_edges_forward(const float* flow, float flow_scale, const float* const* predictions,
                                     const float* const* edges, float lambda_w, const int* out_hw, int num_scales,
                                     int N, int in_h, int in_w, int C, float* const* downsampled,
                                     float* const* residual_dir, float* epe_sums, void* workspace,
                                     size_t workspace_bytes, void* stream);

/* 1 if fn2_multiscale_epe_forward takes these shapes, 0 if it would answer FN2_E_UNSUPPORTED_GEOMETRY (limits above),
 * or a negative fn2_status for arguments the downsample op itself rejects.  Host-only: no CUDA call is made. */
FN2_API int fn2_multiscale_epe_supported(const int* out_hw, int num_scales, int N, int in_h, int in_w, int C);

/* Number of kernel launches (library kernels, not memsets) enqueued by this process so far; bench.py reports the
 * difference over its timed region as `gpu_launches`. */
FN2_API unsigned long long fn2_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* FN2OPS_H_ */
