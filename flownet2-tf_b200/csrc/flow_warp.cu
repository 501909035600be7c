// flow_warp.cu -- bilinear backward-warp of an image by a flow field, forward and backward, for sm_100a.
//
// Semantics follow flow_warp.cu.cc:42-94 and flow_warp_grad.cu.cc:30-87 of the reference:
//   x2 = x + flow.x, y2 = y + flow.y; outside [0,W) x [0,H) (or NaN) -> 0; L = int(x2), T = int(y2),
//   R = min(L+1, W-1), B = min(T+1, H-1); weights from alpha = x2-L, beta = y2-T.
// Both kernels are HBM-bound gathers/scatters: one thread per pixel, the four tap weights in registers, flow read
// as one 8-byte load, and (for the narrow images FlowNet2 warps, C <= 4) outputs staged through shared memory so
// the global stores are full 128-bit coalesced lines instead of C strided scalars.
//
// Narrow images, the tap access: the left and right taps of one image row are 2*C contiguous floats starting at a
// 4-byte-aligned address.  Instead of 2*C scalar gathers (each a separate pass through L1 that touches up to 32
// lines per warp) a thread reads the two or three ALIGNED 16-byte chunks that cover the span (LDG.128) and picks its
// floats out of registers by the span's offset in the first chunk.  The backward scatters the same way: the 2*C
// products are placed into chunk registers (zeros elsewhere, adding 0 is harmless) and sent as 16-byte vector
// reductions (red.global.add.v4.f32), 2-3 per image row instead of 2*C scalar atomics.
#include "fn2_common.cuh"

namespace fn2 {
namespace {

constexpr int WARP_PIX = 256;  // pixels per CTA
constexpr unsigned FULL = 0xffffffffu;

struct Taps {
  bool in;
  int L, T, R, B;
  float alpha, beta, x2, y2;
};

__device__ __forceinline__ Taps make_taps(float fx, float fy, int x, int y, int W, int H) {
  Taps t;
  const float x2 = float(x) + fx, y2 = float(y) + fy;
  t.x2 = x2; t.y2 = y2;
  t.in = (x2 >= 0.f) && (y2 >= 0.f) && (x2 < float(W)) && (y2 < float(H));
  t.L = t.in ? int(x2) : 0;
  t.T = t.in ? int(y2) : 0;
  t.R = min(t.L + 1, W - 1);
  t.B = min(t.T + 1, H - 1);
  t.alpha = x2 - float(t.L);
  t.beta = y2 - float(t.T);
  return t;
}


// ---- aligned-chunk access to the 2*C floats [s, s + 2C) of a row pair (s = absolute float index in the tensor) ----
template <int C>
struct PairSpan {
  static constexpr int NCH = (3 + 2 * C + 3) / 4;   // chunks that can be touched: offset <= 3
};

// value k (0 <= k < 4*NCH) of the chunk registers shifted by off: r[k + off]
template <int NR>
__device__ __forceinline__ float pick(const float (&r)[NR], int k, int off) {
  const float a = r[k], b = (k + 1 < NR) ? r[k + 1] : 0.f, c = (k + 2 < NR) ? r[k + 2] : 0.f,
              d = (k + 3 < NR) ? r[k + 3] : 0.f;
  return off == 0 ? a : (off == 1 ? b : (off == 2 ? c : d));
}

template <int C>
__device__ __forceinline__ void load_pair(const float* __restrict__ t, long long s, long long total, int len,
                                          float (&v)[2 * C]) {
  constexpr int NCH = PairSpan<C>::NCH;
  const long long q = s >> 2;
  const int off = (int)(s & 3);
  float r[4 * NCH];
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const long long f0 = (q + i) * 4;
    float4 c4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (4 * i < off + len) {
      if (f0 + 3 < total) {
        c4 = __ldg(reinterpret_cast<const float4*>(t) + q + i);
      } else {                                              // tail of a tensor whose size is not a multiple of 4
        if (f0 < total) c4.x = __ldg(t + f0);
        if (f0 + 1 < total) c4.y = __ldg(t + f0 + 1);
        if (f0 + 2 < total) c4.z = __ldg(t + f0 + 2);
      }
    }
    r[4 * i] = c4.x; r[4 * i + 1] = c4.y; r[4 * i + 2] = c4.z; r[4 * i + 3] = c4.w;
  }
#pragma unroll
  for (int k = 0; k < 2 * C; ++k) v[k] = pick(r, k, off);
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// adds v[0 .. len) to t[s .. s + len): aligned 16-byte vector reductions, zeros in the lanes outside the span
template <int C>
__device__ __forceinline__ void scatter_pair(float* __restrict__ t, long long s, long long total, int len,
                                             const float (&v)[2 * C]) {
  constexpr int NCH = PairSpan<C>::NCH;
  const long long q = s >> 2;
  const int off = (int)(s & 3);
  float r[4 * NCH];
#pragma unroll
  for (int k = 0; k < 4 * NCH; ++k) {
    // r[k] = v[k - off] when that index is inside [0, len)
    const float a = (k < 2 * C) ? v[k] : 0.f;
    const float b = (k >= 1 && k - 1 < 2 * C) ? v[k - 1] : 0.f;
    const float c = (k >= 2 && k - 2 < 2 * C) ? v[k - 2] : 0.f;
    const float d = (k >= 3 && k - 3 < 2 * C) ? v[k - 3] : 0.f;
    const float x = off == 0 ? a : (off == 1 ? b : (off == 2 ? c : d));
    r[k] = (k - off < len) ? x : 0.f;
  }
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    if (4 * i < off + len) {
      const long long f0 = (q + i) * 4;
      if (f0 + 3 < total) {
        red_add_v4(t + f0, r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]);
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (f0 + e < total && r[4 * i + e] != 0.f) atomicAdd(t + f0 + e, r[4 * i + e]);
      }
    }
  }
}

// ---- forward, narrow images: thread per pixel, smem-staged coalesced stores -------------------------------
template <int C>
__global__ void __launch_bounds__(WARP_PIX) warp_fwd_narrow(const float* __restrict__ image,
                                                            const float2* __restrict__ flow,
                                                            float* __restrict__ out, long long npix, int H, int W) {
  __shared__ __align__(16) float stage[WARP_PIX * C];
  const long long base = (long long)blockIdx.x * WARP_PIX;
  const long long pix = base + threadIdx.x;
  float v[C];
#pragma unroll
  for (int c = 0; c < C; ++c) v[c] = 0.f;
  if (pix < npix) {
    const int x = (int)(pix % W);
    const int y = (int)((pix / W) % H);
    const long long n = pix / ((long long)W * H);
    const float2 f = __ldg(flow + pix);
    const Taps t = make_taps(f.x, f.y, x, y, W, H);
    if (t.in) {
      const long long total = npix * C;
      const int len = (t.R == t.L) ? C : 2 * C;            // right tap clamped onto the left one (:56)
      float top[2 * C], bot[2 * C];
      load_pair<C>(image, ((n * H + t.T) * (long long)W + t.L) * C, total, len, top);
      load_pair<C>(image, ((n * H + t.B) * (long long)W + t.L) * C, total, len, bot);
      const float cTL = (1 - t.alpha) * (1 - t.beta), cTR = t.alpha * (1 - t.beta);
      const float cBL = (1 - t.alpha) * t.beta, cBR = t.alpha * t.beta;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const float iTR = (t.R == t.L) ? top[c] : top[C + c], iBR = (t.R == t.L) ? bot[c] : bot[C + c];
        v[c] = cTL * top[c] + cTR * iTR + cBL * bot[c] + cBR * iBR;
      }
    }
  }
#pragma unroll
  for (int c = 0; c < C; ++c) stage[threadIdx.x * C + c] = v[c];
  __syncthreads();
  // WARP_PIX*C floats, 16-byte aligned in global because WARP_PIX*C*4 is a multiple of 16
  const long long remaining = (npix - base) * C;  // floats left from this CTA's base
  const int nfl = (int)min((long long)WARP_PIX * C, remaining);
  float* dst = out + (size_t)base * C;
  const int nvec = nfl >> 2;
  for (int i = threadIdx.x; i < nvec; i += WARP_PIX)
    reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(stage)[i];
  for (int i = (nvec << 2) + threadIdx.x; i < nfl; i += WARP_PIX) dst[i] = stage[i];
}

// ---- forward, wide tensors: one warp per pixel, lanes over channels ---------------------------------------
template <bool VEC4>
__global__ void __launch_bounds__(256) warp_fwd_wide(const float* __restrict__ image, const float2* __restrict__ flow,
                                                     float* __restrict__ out, long long npix, int H, int W, int C) {
  const int lane = threadIdx.x & 31;
  const long long pix = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (pix >= npix) return;
  const int x = (int)(pix % W);
  const int y = (int)((pix / W) % H);
  const long long n = pix / ((long long)W * H);
  const float2 f = __ldg(flow + pix);
  const Taps t = make_taps(f.x, f.y, x, y, W, H);
  float* po = out + (size_t)pix * C;
  if (!t.in) {
    if (VEC4) for (int c = lane * 4; c < C; c += 128) *reinterpret_cast<float4*>(po + c) = make_float4(0, 0, 0, 0);
    else for (int c = lane; c < C; c += 32) po[c] = 0.f;
    return;
  }
  const float* img = image + (size_t)n * H * W * C;
  const float* pTL = img + ((size_t)t.T * W + t.L) * C;
  const float* pTR = img + ((size_t)t.T * W + t.R) * C;
  const float* pBL = img + ((size_t)t.B * W + t.L) * C;
  const float* pBR = img + ((size_t)t.B * W + t.R) * C;
  const float cTL = (1 - t.alpha) * (1 - t.beta), cTR = t.alpha * (1 - t.beta);
  const float cBL = (1 - t.alpha) * t.beta, cBR = t.alpha * t.beta;
  if (VEC4) {
    for (int c = lane * 4; c < C; c += 128) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(pTL + c)), b = __ldg(reinterpret_cast<const float4*>(pTR + c));
      const float4 d = __ldg(reinterpret_cast<const float4*>(pBL + c)), e = __ldg(reinterpret_cast<const float4*>(pBR + c));
      float4 r;
      r.x = cTL * a.x + cTR * b.x + cBL * d.x + cBR * e.x;
      r.y = cTL * a.y + cTR * b.y + cBL * d.y + cBR * e.y;
      r.z = cTL * a.z + cTR * b.z + cBL * d.z + cBR * e.z;
      r.w = cTL * a.w + cTR * b.w + cBL * d.w + cBR * e.w;
      *reinterpret_cast<float4*>(po + c) = r;
    }
  } else {
    for (int c = lane; c < C; c += 32)
      po[c] = cTL * __ldg(pTL + c) + cTR * __ldg(pTR + c) + cBL * __ldg(pBL + c) + cBR * __ldg(pBR + c);
  }
}

// ---- backward, narrow images -------------------------------------------------------------------------------
// One thread per pixel, one warp = 32 consecutive pixels of the flattened image.  With a locally smooth flow the
// right taps of pixel x are the left taps of pixel x+1, so before touching memory each lane hands its right-column
// contributions to its neighbour when the targets coincide ("warp-aggregated" atomics): 2 instead of 4 scatter
// targets per pixel in the common case.  Remaining adds are fire-and-forget red.global.add.f32.
template <int C>
__global__ void __launch_bounds__(WARP_PIX) warp_bwd_narrow(const float* __restrict__ image,
                                                            const float2* __restrict__ flow,
                                                            const float* __restrict__ gout,
                                                            float* __restrict__ gimage, float2* __restrict__ gflow,
                                                            long long npix, int H, int W) {
  __shared__ __align__(16) float stage[WARP_PIX * C];
  const long long base = (long long)blockIdx.x * WARP_PIX;
  const long long pix = base + threadIdx.x;
  const int lane = threadIdx.x & 31;

  // coalesced 128-bit read of this CTA's gradients into shared memory
  {
    const long long remaining = (npix - base) * C;
    const int nfl = (int)min((long long)WARP_PIX * C, remaining);
    const float* src = gout + (size_t)base * C;
    const int nvec = nfl >> 2;
    for (int i = threadIdx.x; i < nvec; i += WARP_PIX)
      reinterpret_cast<float4*>(stage)[i] = __ldg(reinterpret_cast<const float4*>(src) + i);
    for (int i = (nvec << 2) + threadIdx.x; i < nfl; i += WARP_PIX) stage[i] = __ldg(src + i);
  }
  __syncthreads();

  const bool live = pix < npix;
  Taps t;
  t.in = false; t.L = t.T = t.R = t.B = 0; t.alpha = t.beta = t.x2 = t.y2 = 0.f;
  long long n = 0;
  float g[C];
#pragma unroll
  for (int c = 0; c < C; ++c) g[c] = 0.f;
  if (live) {
    const int x = (int)(pix % W);
    const int y = (int)((pix / W) % H);
    n = pix / ((long long)W * H);
    const float2 f = __ldg(flow + pix);
    t = make_taps(f.x, f.y, x, y, W, H);
#pragma unroll
    for (int c = 0; c < C; ++c) g[c] = stage[threadIdx.x * C + c];
  }
  const bool act = live && t.in;

  float wTL[C], wTR[C], wBL[C], wBR[C];
#pragma unroll
  for (int c = 0; c < C; ++c) {
    wTL[c] = g[c] * (1 - t.alpha) * (1 - t.beta);
    wTR[c] = g[c] * t.alpha * (1 - t.beta);
    wBL[c] = g[c] * (1 - t.alpha) * t.beta;
    wBR[c] = g[c] * t.alpha * t.beta;
  }

  // ---- d/dflow (needs the four taps of the image) ----
  float2 gf = make_float2(0.f, 0.f);
  const long long total = npix * C;
  const long long sT = ((n * H + t.T) * (long long)W + t.L) * C, sB = ((n * H + t.B) * (long long)W + t.L) * C;
  const bool clampR = (t.R == t.L);
  if (act) {
    const float gy = float(t.B) - t.y2;  // flow_warp_grad.cu.cc:55
    const float gx = float(t.R) - t.x2;  // :72
    float top[2 * C], bot[2 * C];
    load_pair<C>(image, sT, total, clampR ? C : 2 * C, top);
    load_pair<C>(image, sB, total, clampR ? C : 2 * C, bot);
    float bx = 0.f, by = 0.f;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const float iTL = top[c], iTR = clampR ? top[c] : top[C + c];
      const float iBL = bot[c], iBR = clampR ? bot[c] : bot[C + c];
      float tx = gy * (iTR - iTL);
      tx += (1 - gy) * (iBR - iBL);
      bx += g[c] * tx;
      float ty = gx * (iBL - iTL);
      ty += (1 - gx) * (iBR - iTR);
      by += g[c] * ty;
    }
    gf = make_float2(bx, by);
  }
  if (live) gflow[pix] = gf;  // 8-byte coalesced store

  // ---- d/dimage: neighbour aggregation, then vector reductions ----
  // key of my left column / right column targets (same image n within a warp unless the warp straddles images)
  const long long keyL = act ? ((n * H + t.T) * (long long)W + t.L) : -1;
  const long long keyR = act ? ((n * H + t.T) * (long long)W + t.R) : -2;
  const int myB = t.B;
  // what the lane to my left would send to its right column
  const long long upKeyR = __shfl_up_sync(FULL, keyR, 1);
  const int upB = __shfl_up_sync(FULL, myB, 1);
  const bool absorb = (lane > 0) && act && !clampR && (upKeyR == keyL) && (upB == myB);
  // does the lane to my right absorb my right column?
  const bool given = __shfl_down_sync(FULL, absorb ? 1 : 0, 1) && (lane < 31);
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const float upTR = __shfl_up_sync(FULL, wTR[c], 1);
    const float upBR = __shfl_up_sync(FULL, wBR[c], 1);
    if (absorb) { wTL[c] += upTR; wBL[c] += upBR; }
  }
  if (act) {
    // right taps that coincide with the left ones (clamped) or that the neighbour took over leave the span
    const int len = (clampR || given) ? C : 2 * C;
    float vt[2 * C], vb[2 * C];
#pragma unroll
    for (int c = 0; c < C; ++c) {
      vt[c] = wTL[c] + ((clampR && !given) ? wTR[c] : 0.f);
      vb[c] = wBL[c] + ((clampR && !given) ? wBR[c] : 0.f);
      vt[C + c] = wTR[c];
      vb[C + c] = wBR[c];
    }
    if (t.B == t.T) {                                      // bottom taps clamped onto the top row (:57)
#pragma unroll
      for (int k = 0; k < 2 * C; ++k) vt[k] += vb[k];
      scatter_pair<C>(gimage, sT, total, len, vt);
    } else {
      scatter_pair<C>(gimage, sT, total, len, vt);
      scatter_pair<C>(gimage, sB, total, len, vb);
    }
  }
}

// ---- backward, wide tensors: warp per pixel, lanes over channels -------------------------------------------
__global__ void __launch_bounds__(256) warp_bwd_wide(const float* __restrict__ image, const float2* __restrict__ flow,
                                                     const float* __restrict__ gout, float* __restrict__ gimage,
                                                     float2* __restrict__ gflow, long long npix, int H, int W, int C) {
  const int lane = threadIdx.x & 31;
  const long long pix = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (pix >= npix) return;
  const int x = (int)(pix % W);
  const int y = (int)((pix / W) % H);
  const long long n = pix / ((long long)W * H);
  const float2 f = __ldg(flow + pix);
  const Taps t = make_taps(f.x, f.y, x, y, W, H);
  if (!t.in) {
    if (lane == 0) gflow[pix] = make_float2(0.f, 0.f);
    return;
  }
  const float* img = image + (size_t)n * H * W * C;
  float* gi = gimage + (size_t)n * H * W * C;
  const size_t oTL = ((size_t)t.T * W + t.L) * C, oTR = ((size_t)t.T * W + t.R) * C;
  const size_t oBL = ((size_t)t.B * W + t.L) * C, oBR = ((size_t)t.B * W + t.R) * C;
  const float gy = float(t.B) - t.y2, gx = float(t.R) - t.x2;
  const float cTL = (1 - t.alpha) * (1 - t.beta), cTR = t.alpha * (1 - t.beta);
  const float cBL = (1 - t.alpha) * t.beta, cBR = t.alpha * t.beta;
  float bx = 0.f, by = 0.f;
  for (int c = lane; c < C; c += 32) {
    const float gv = __ldg(gout + (size_t)pix * C + c);
    atomicAdd(gi + oTL + c, gv * cTL);
    atomicAdd(gi + oTR + c, gv * cTR);
    atomicAdd(gi + oBL + c, gv * cBL);
    atomicAdd(gi + oBR + c, gv * cBR);
    const float iTL = __ldg(img + oTL + c), iTR = __ldg(img + oTR + c);
    const float iBL = __ldg(img + oBL + c), iBR = __ldg(img + oBR + c);
    float tx = gy * (iTR - iTL);
    tx += (1 - gy) * (iBR - iBL);
    bx += gv * tx;
    float ty = gx * (iBL - iTL);
    ty += (1 - gx) * (iBR - iTR);
    by += gv * ty;
  }
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    bx += __shfl_xor_sync(FULL, bx, off);
    by += __shfl_xor_sync(FULL, by, off);
  }
  if (lane == 0) gflow[pix] = make_float2(bx, by);
}

template <int C>
int launch_fwd_narrow(const float* image, const float* flow, float* out, long long npix, int H, int W,
                      cudaStream_t st) {
  const long long blocks = (npix + WARP_PIX - 1) / WARP_PIX;
  warp_fwd_narrow<C><<<(unsigned)blocks, WARP_PIX, 0, st>>>(image, reinterpret_cast<const float2*>(flow), out, npix,
                                                             H, W);
  return after_launch();
}
template <int C>
int launch_bwd_narrow(const float* image, const float* flow, const float* gout, float* gimage, float* gflow,
                      long long npix, int H, int W, cudaStream_t st) {
  const long long blocks = (npix + WARP_PIX - 1) / WARP_PIX;
  warp_bwd_narrow<C><<<(unsigned)blocks, WARP_PIX, 0, st>>>(image, reinterpret_cast<const float2*>(flow), gout,
                                                             gimage, reinterpret_cast<float2*>(gflow), npix, H, W);
  return after_launch();
}

}  // namespace

int flow_warp_forward(const float* image, const float* flow, float* out, int N, int H, int W, int C,
                      cudaStream_t st) {
  const long long npix = (long long)N * H * W;
  if ((npix + WARP_PIX - 1) / WARP_PIX > 0x7fffffffLL || (npix + 7) / 8 > 0x7fffffffLL) return FN2_E_BAD_SHAPE;
  if (flow_warp_tile_supported(H, W, C)) return flow_warp_tile_forward(image, flow, out, N, H, W, C, st);
  switch (C) {
    case 1: return launch_fwd_narrow<1>(image, flow, out, npix, H, W, st);
    case 2: return launch_fwd_narrow<2>(image, flow, out, npix, H, W, st);
    case 3: return launch_fwd_narrow<3>(image, flow, out, npix, H, W, st);
    case 4: return launch_fwd_narrow<4>(image, flow, out, npix, H, W, st);
    default: break;
  }
  const long long blocks = (npix + 7) / 8;
  if (C % 4 == 0)
    warp_fwd_wide<true><<<(unsigned)blocks, 256, 0, st>>>(image, reinterpret_cast<const float2*>(flow), out, npix, H, W, C);
  else
    warp_fwd_wide<false><<<(unsigned)blocks, 256, 0, st>>>(image, reinterpret_cast<const float2*>(flow), out, npix, H, W, C);
  return after_launch();
}

namespace {
// err[pixel] = sqrt(sum_c (reference - warped)^2): the shapes the window kernel does not take
__global__ void __launch_bounds__(256) channel_diff_norm_kernel(const float* __restrict__ ref,
                                                                const float* __restrict__ warped,
                                                                float* __restrict__ err, long long npix, int C) {
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += (long long)gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int c = 0; c < C; ++c) {
      const float d = __ldg(ref + p * C + c) - warped[p * C + c];
      acc += d * d;
    }
    err[p] = sqrtf(acc);
  }
}
}  // namespace

int flow_warp_forward_error(const float* image, const float* flow, const float* reference, float* out, float* err, int N,
                            int H, int W, int C, cudaStream_t st) {
  if (flow_warp_tile_supported(H, W, C)) return flow_warp_tile_forward_error(image, flow, reference, out, err, N, H, W, C, st);
  const int rc = flow_warp_forward(image, flow, out, N, H, W, C, st);
  if (rc != FN2_OK) return rc;
  const long long npix = (long long)N * H * W;
  long long blocks = (npix + 255) / 256;
  if (blocks > 148LL * 16) blocks = 148LL * 16;
  channel_diff_norm_kernel<<<(unsigned)blocks, 256, 0, st>>>(reference, out, err, npix, C);
  return after_launch();
}

int flow_warp_backward(const float* image, const float* flow, const float* gout, float* gimage, float* gflow, int N,
                       int H, int W, int C, cudaStream_t st) {
  const long long npix = (long long)N * H * W;
  if ((npix + WARP_PIX - 1) / WARP_PIX > 0x7fffffffLL || (npix + 7) / 8 > 0x7fffffffLL) return FN2_E_BAD_SHAPE;
  cudaError_t e = cudaMemsetAsync(gimage, 0, sizeof(float) * (size_t)npix * C, st);
  if (e != cudaSuccess) { note_cuda_error(e); return FN2_E_CUDA; }
  if (flow_warp_tile_supported(H, W, C)) return flow_warp_tile_backward(image, flow, gout, gimage, gflow, N, H, W, C, st);
  switch (C) {
    case 1: return launch_bwd_narrow<1>(image, flow, gout, gimage, gflow, npix, H, W, st);
    case 2: return launch_bwd_narrow<2>(image, flow, gout, gimage, gflow, npix, H, W, st);
    case 3: return launch_bwd_narrow<3>(image, flow, gout, gimage, gflow, npix, H, W, st);
    case 4: return launch_bwd_narrow<4>(image, flow, gout, gimage, gflow, npix, H, W, st);
    default: break;
  }
  const long long blocks = (npix + 7) / 8;
  warp_bwd_wide<<<(unsigned)blocks, 256, 0, st>>>(image, reinterpret_cast<const float2*>(flow), gout, gimage,
                                                   reinterpret_cast<float2*>(gflow), npix, H, W, C);
  return after_launch();
}

}  // namespace fn2
