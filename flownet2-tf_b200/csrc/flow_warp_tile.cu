// flow_warp_tile.cu -- flow_warp forward / backward for narrow images (C <= 4) with the sampled neighbourhood staged
// in shared memory.  Semantics: flow_warp.cu.cc:42-94, flow_warp_grad.cu.cc:30-87.
//
// A per-pixel gather straight from global memory is bound by L1 wavefronts: with flows of a few pixels every lane of
// a warp touches its own pair of 32-byte sectors in two image rows.  But the taps of a TILE of output pixels all fall
// in the tile grown by the flow magnitude.  One CTA owns a 32 x 64 tile of output pixels, has the TMA unit copy the
// window [tile - margin, tile + margin] of the image into shared memory (cp.async.bulk.tensor: no LSU instructions,
// parts of the window outside the image are zero fill; each image byte is fetched from L2 once per neighbouring tile
// instead of once per tap), and the four taps become shared-memory reads.  A pixel whose taps leave the window (flow larger than the margin) takes them from global memory instead:
// any flow is handled, only the common case is fast.
//
// Backward, image gradient: the same window, zero-initialised, collects the four-tap scatter with shared-memory
// adds; it is then flushed row by row with bulk reduce-adds (cp.reduce.async.bulk .add.f32) instead of 12 scattered
// atomics per pixel.  Out-of-window taps go to global atomics directly.
// Backward, flow gradient: a gather like the forward, from the staged image window.
#include <cstdlib>
#include <mutex>
#include "fn2_tc.cuh"

namespace fn2 {
namespace {

constexpr int TH = 32, TW = 64;        // output tile
constexpr int NT = 512;                // threads per CTA: 4 consecutive pixels each
constexpr int PXT = 4;

using namespace tc;

template <int C>
struct Win {
  // C <= 3: margins 12 x 12 -> 57 x 96 pixels = 66 KB (three CTAs per SM); C == 4: 16 x 16 -> 65 x 104 pixels
#ifndef FN2_WIN_MY   // developer A/B of the gather window (variant builds, tools/gpu_ab.sh)
#define FN2_WIN_MY 12   // measured (tools/time_warp.py, N=32; forward / backward ms): margins 22x24 0.0589 / 0.218 on
#define FN2_WIN_MX 12   // N(0, 8 px) flows and 0.0675 / 0.209 on the reference's sample flow; 16x16 0.0576 / 0.208 and
#define FN2_WIN_COLS 96 // 0.0654 / 0.201; 12x12 0.0583 / 0.2105 and 0.0607 / 0.195; 8x8 0.0613 / 0.216 and 0.0628 / 0.194
#endif
  static constexpr int MY = (C <= 3) ? FN2_WIN_MY : 16;
  static constexpr int MX = (C <= 3) ? FN2_WIN_MX : 16;
  static constexpr int ROWS = TH + 2 * MY + 1;
  static constexpr int COLS = (C <= 3) ? FN2_WIN_COLS : 104;                  // >= TW + 2*MX + 1
  static constexpr int NB = (COLS * C + 255) / 256;                  // TMA boxes (inner box dimension <= 256 elements)
  static constexpr int PXB = COLS / NB;                              // pixels per box
  static constexpr int BW = PXB * C;                                 // floats per box row
  static constexpr int BOX_FLOATS = ((ROWS * BW * 4 + 127) / 128) * 32;   // box size rounded up to 128 bytes
  static constexpr uint32_t TX_BYTES = (uint32_t)NB * ROWS * BW * 4;
  static constexpr size_t BYTES = (size_t)NB * BOX_FLOATS * sizeof(float) + 128;
  static_assert(COLS % NB == 0 && (BW * 4) % 16 == 0 && COLS >= TW + 2 * MX + 1, "window geometry");
  // float offset of pixel column px (0 <= px < COLS) of window row r
  __device__ static __forceinline__ int at(int r, int px) {
    const int box = (NB > 1 && px >= PXB) ? 1 : 0;
    return box * BOX_FLOATS + r * BW + (px - box * PXB) * C;
  }
};

struct Tile {
  int n, ty0, tx0;          // image, tile origin
  int wy0, wx0;             // window origin (may be negative: zero fill)
};

template <int C>
__device__ __forceinline__ Tile make_tile(int tilesX, int tilesY) {
  Tile t;
  int b = blockIdx.x;
  const int txi = b % tilesX; b /= tilesX;
  const int tyi = b % tilesY;
  t.n = b / tilesY;
  t.ty0 = tyi * TH; t.tx0 = txi * TW;
  t.wy0 = t.ty0 - Win<C>::MY;
  t.wx0 = t.tx0 - Win<C>::MX;
  return t;
}

__device__ __forceinline__ float* align128(float* p) {
  return p + (((128u - (smem_u32(p) & 127u)) & 127u) >> 2);
}

struct Taps {
  bool in;
  int L, T, R, B;
  float alpha, beta, x2, y2;
};

__device__ __forceinline__ Taps make_taps(float fx, float fy, int x, int y, int W, int H) {
  Taps t;
  const float x2 = float(x) + fx, y2 = float(y) + fy;
  t.x2 = x2; t.y2 = y2;
  t.in = (x2 >= 0.f) && (y2 >= 0.f) && (x2 < float(W)) && (y2 < float(H));   // NaN -> false (flow_warp.cu.cc:74)
  t.L = t.in ? int(x2) : 0;
  t.T = t.in ? int(y2) : 0;
  t.R = min(t.L + 1, W - 1);
  t.B = min(t.T + 1, H - 1);
  t.alpha = x2 - float(t.L);
  t.beta = y2 - float(t.T);
  return t;
}

// Image-gradient scatter window: a smaller margin than the gather window.  Scattered atomics cost about one lane per
// clock whatever unit executes them (shared-memory fp32 adds are LDS + ATOMS.CAST.SPIN, two lane operations per value);
// a tap outside the window costs FEWER lane operations as 8-byte global reductions (red.global.add.v2.f32, two floats
// per lane operation) than one inside, and a smaller window is cheaper to zero and to flush, so the window only has to
// catch the taps whose flush would otherwise be a scattered reduction per float: measured (tools/mb/mb_scatter.cu,
// batch 32, flows ~ N(0, 8 px)): margins 22x24 0.195 ms, 16x16 0.164, 12x12 0.152, 8x8 0.142, 4x4 0.137; the same
// ranking for sigma 2 and 0.5.
template <int C>
struct Scat {
  static constexpr int MY = 8, MX = 8;                              // MX % 4 == 0: window rows start 16-byte aligned
  static constexpr int ROWS = TH + 2 * MY + 1;
  static constexpr int COLS = (TW + 2 * MX + 1 + 3) / 4 * 4;
  static constexpr int RF = COLS * C;                               // floats per window row
  static constexpr int FLOATS = ROWS * RF;
  static_assert(FLOATS * 4 <= (int)Win<C>::BYTES - 128, "the scatter window lives in the gather window's buffer");
};

// adds v[0 .. 2C) to the 2C consecutive floats at p (left and right tap of one image row): 8-byte reductions for the
// aligned pairs, scalar ones for a ragged first / last float
template <int C>
__device__ __forceinline__ void red_run(float* p, const float (&v)[2 * C]) {
  if ((reinterpret_cast<uintptr_t>(p) & 7u) == 0) {
#pragma unroll
    for (int k = 0; k < C; ++k) atomicAdd(reinterpret_cast<float2*>(p) + k, make_float2(v[2 * k], v[2 * k + 1]));
  } else {
    atomicAdd(p, v[0]);
#pragma unroll
    for (int k = 0; k < C - 1; ++k)
      atomicAdd(reinterpret_cast<float2*>(p + 1) + k, make_float2(v[2 * k + 1], v[2 * k + 2]));
    atomicAdd(p + 2 * C - 1, v[2 * C - 1]);
  }
}

// The four-tap scatter of one pixel's incoming gradient (flow_warp_grad.cu.cc:43-53): into the zeroed shared-memory
// window `win` (origin wy0, wx0) when all four taps fall inside it, else straight to global memory.
template <int C>
__device__ __forceinline__ void scatter_taps(const Taps& tp, const float* gv, float* win, int wy0, int wx0,
                                             float* gimg, int W) {
  const float cTL = (1 - tp.alpha) * (1 - tp.beta), cTR = tp.alpha * (1 - tp.beta);
  const float cBL = (1 - tp.alpha) * tp.beta, cBR = tp.alpha * tp.beta;
  const bool inwin = (tp.T >= wy0) && (tp.B < wy0 + Scat<C>::ROWS) && (tp.L >= wx0) && (tp.R < wx0 + Scat<C>::COLS);
  if (inwin) {
    float* pL = win + (tp.T - wy0) * Scat<C>::RF + (tp.L - wx0) * C;
    float* pR = win + (tp.T - wy0) * Scat<C>::RF + (tp.R - wx0) * C;
    const int dB = (tp.B - tp.T) * Scat<C>::RF;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const float gc = gv[c];
      atomicAdd(pL + c, gc * cTL);
      atomicAdd(pR + c, gc * cTR);
      atomicAdd(pL + dB + c, gc * cBL);
      atomicAdd(pR + dB + c, gc * cBR);
    }
  } else if (tp.R == tp.L + 1) {
    // the left and right tap of a row are 2C consecutive floats
    float vt[2 * C], vb[2 * C];
#pragma unroll
    for (int c = 0; c < C; ++c) {
      vt[c] = gv[c] * cTL; vt[C + c] = gv[c] * cTR;
      vb[c] = gv[c] * cBL; vb[C + c] = gv[c] * cBR;
    }
    float* pT = gimg + ((size_t)tp.T * W + tp.L) * C;
    red_run<C>(pT, vt);
    red_run<C>(pT + (size_t)(tp.B - tp.T) * W * C, vb);    // B == T (last image row): the same floats again, as the
  } else {                                                  // reference's two atomics to one address
    float* pT = gimg + ((size_t)tp.T * W + tp.L) * C;      // R == L: last image column
    const size_t dB = (size_t)(tp.B - tp.T) * W * C;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const float gc = gv[c];
      atomicAdd(pT + c, gc * cTL);
      atomicAdd(pT + c, gc * cTR);
      atomicAdd(pT + dB + c, gc * cBL);
      atomicAdd(pT + dB + c, gc * cBR);
    }
  }
}

// zero the scatter window / flush it: every window row inside the image is one bulk reduce-add
// (cp.reduce.async.bulk .add.f32, executed by the copy engine: no LSU instructions), clipped to the image; offsets and
// sizes stay multiples of 16 bytes because W*C, MX*C and the window width are multiples of 4 floats.
template <int C>
__device__ __forceinline__ void scatter_zero(float* win) {
  for (int i = threadIdx.x; i < Scat<C>::FLOATS / 4; i += NT) reinterpret_cast<float4*>(win)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}
template <int C>
__device__ __forceinline__ void scatter_flush(const float* win, float* gimage, int n, int H, int W, int wy0, int wx0) {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the async proxy
  __syncthreads();
  if (threadIdx.x < Scat<C>::ROWS) {
    const int r = threadIdx.x;
    const int yy = wy0 + r;
    const int fc0 = wx0 * C;
    const int lo = max(fc0, 0), hi = min(fc0 + Scat<C>::RF, W * C);
    if (yy >= 0 && yy < H && hi > lo) {
      float* dst = gimage + ((size_t)n * H + yy) * W * C + lo;
      const uint32_t src = smem_u32(win + r * Scat<C>::RF + (lo - fc0));
      asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
                   ::"l"(dst), "r"(src), "r"((uint32_t)(hi - lo) * 4u) : "memory");
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // shared memory stays valid until it was read
  }
}

// ------------------------------------------------------------------------------------------------ forward
// MODE 0: out = warp(image, flow).   MODE 1: flow gradient (needs gout), same gather.
// MODE 2: MODE 0 plus the brightness error of the stacked nets (flownet_cs.py:24-27, flownet2.py:33-39): gout is
//         the reference image (input_a) and out2[pixel] = sqrt(sum_c (reference - warped)^2), in the same pass.
template <int C, int MODE>
__global__ void __launch_bounds__(NT, 2) warp_tile_gather(const __grid_constant__ CUtensorMap imap,
                                                          const float* __restrict__ image,
                                                          const float* __restrict__ flow,
                                                          const float* __restrict__ gout, float* __restrict__ out,
                                                          float* __restrict__ out2, int H, int W, int tilesX,
                                                          int tilesY) {
  extern __shared__ __align__(128) float win_raw[];
  __shared__ __align__(8) uint64_t bar;
  float* win = align128(win_raw);
  const Tile t = make_tile<C>(tilesX, tilesY);
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    mbar_expect_tx(smem_u32(&bar), Win<C>::TX_BYTES);
#pragma unroll
    for (int b = 0; b < Win<C>::NB; ++b)
      tma_load_3d(smem_u32(win + b * Win<C>::BOX_FLOATS), &imap, smem_u32(&bar), t.wx0 * C + b * Win<C>::BW, t.wy0,
                  t.n);
  }
  __syncthreads();   // the barrier is initialised before anyone polls it
  const int y = t.ty0 + threadIdx.x / (TW / PXT);
  const int x = t.tx0 + (threadIdx.x % (TW / PXT)) * PXT;
  const bool live = (y < H) && (x < W);          // W % 4 == 0: a group of 4 pixels is inside or outside as a whole
  const size_t pix = ((size_t)t.n * H + (live ? y : 0)) * W + (live ? x : 0);

  float2 f[PXT];
  float g[PXT][C];
  if (live) {   // issued before the window copy so the latencies overlap
    const float4 f01 = __ldg(reinterpret_cast<const float4*>(flow + pix * 2));
    const float4 f23 = __ldg(reinterpret_cast<const float4*>(flow + pix * 2) + 1);
    f[0] = make_float2(f01.x, f01.y); f[1] = make_float2(f01.z, f01.w);
    f[2] = make_float2(f23.x, f23.y); f[3] = make_float2(f23.z, f23.w);
    if (MODE >= 1) {
      float gv[PXT * C];
#pragma unroll
      for (int v = 0; v < C; ++v) {
        const float4 q = __ldg(reinterpret_cast<const float4*>(gout + pix * C) + v);
        gv[4 * v] = q.x; gv[4 * v + 1] = q.y; gv[4 * v + 2] = q.z; gv[4 * v + 3] = q.w;
      }
#pragma unroll
      for (int i = 0; i < PXT; ++i)
#pragma unroll
        for (int c = 0; c < C; ++c) g[i][c] = gv[i * C + c];
    }
  }
  mbar_wait(smem_u32(&bar), 0);
  if (!live) return;

  float res[PXT * C];     // MODE 0: warped pixels
  float2 gf[PXT];         // MODE 1: flow gradients
  const float* img = image + (size_t)t.n * H * W * C;
#pragma unroll
  for (int i = 0; i < PXT; ++i) {
    const Taps tp = make_taps(f[i].x, f[i].y, x + i, y, W, H);
    float iTL[C], iTR[C], iBL[C], iBR[C];
#pragma unroll
    for (int c = 0; c < C; ++c) iTL[c] = iTR[c] = iBL[c] = iBR[c] = 0.f;
    if (tp.in) {
      const bool inwin = (tp.T >= t.wy0) && (tp.B < t.wy0 + Win<C>::ROWS) && (tp.L >= t.wx0) &&
                         (tp.R < t.wx0 + Win<C>::COLS);
      if (inwin) {
        const float* pL = win + Win<C>::at(tp.T - t.wy0, tp.L - t.wx0);
        const float* pR = win + Win<C>::at(tp.T - t.wy0, tp.R - t.wx0);
        const int dB = (tp.B - tp.T) * Win<C>::BW;
#pragma unroll
        for (int c = 0; c < C; ++c) {
          iTL[c] = pL[c]; iTR[c] = pR[c]; iBL[c] = pL[dB + c]; iBR[c] = pR[dB + c];
        }
      } else {
        const float* pT = img + ((size_t)tp.T * W + tp.L) * C;
        const size_t dR = (size_t)(tp.R - tp.L) * C, dB = (size_t)(tp.B - tp.T) * W * C;
#pragma unroll
        for (int c = 0; c < C; ++c) {
          iTL[c] = __ldg(pT + c); iTR[c] = __ldg(pT + dR + c);
          iBL[c] = __ldg(pT + dB + c); iBR[c] = __ldg(pT + dB + dR + c);
        }
      }
    }
    if (MODE != 1) {
      const float cTL = (1 - tp.alpha) * (1 - tp.beta), cTR = tp.alpha * (1 - tp.beta);
      const float cBL = (1 - tp.alpha) * tp.beta, cBR = tp.alpha * tp.beta;
#pragma unroll
      for (int c = 0; c < C; ++c)
        res[i * C + c] = tp.in ? cTL * iTL[c] + cTR * iTR[c] + cBL * iBL[c] + cBR * iBR[c] : 0.f;
    } else {
      const float gy = float(tp.B) - tp.y2;  // flow_warp_grad.cu.cc:55
      const float gx = float(tp.R) - tp.x2;  // :72
      float bx = 0.f, by = 0.f;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        float tx = gy * (iTR[c] - iTL[c]);
        tx += (1 - gy) * (iBR[c] - iBL[c]);
        bx += g[i][c] * tx;
        float ty = gx * (iBL[c] - iTL[c]);
        ty += (1 - gx) * (iBR[c] - iTR[c]);
        by += g[i][c] * ty;
      }
      gf[i] = tp.in ? make_float2(bx, by) : make_float2(0.f, 0.f);
    }
  }
  if (MODE != 1) {
#pragma unroll
    for (int v = 0; v < C; ++v)
      reinterpret_cast<float4*>(out + pix * C)[v] = make_float4(res[4 * v], res[4 * v + 1], res[4 * v + 2], res[4 * v + 3]);
    if (MODE == 2) {
      float e[PXT];
#pragma unroll
      for (int i = 0; i < PXT; ++i) {
        float acc = 0.f;
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const float d = g[i][c] - res[i * C + c];
          acc += d * d;
        }
        e[i] = sqrtf(acc);
      }
      *reinterpret_cast<float4*>(out2 + pix) = make_float4(e[0], e[1], e[2], e[3]);
    }
  } else {
    reinterpret_cast<float4*>(out + pix * 2)[0] = make_float4(gf[0].x, gf[0].y, gf[1].x, gf[1].y);
    reinterpret_cast<float4*>(out + pix * 2)[1] = make_float4(gf[2].x, gf[2].y, gf[3].x, gf[3].y);
  }
}

// ------------------------------------------------------------------------------------- backward: image gradient
template <int C>
__global__ void __launch_bounds__(NT, 2) warp_tile_scatter(const float* __restrict__ flow,
                                                           const float* __restrict__ gout,
                                                           float* __restrict__ gimage, int H, int W, int tilesX,
                                                           int tilesY) {
  extern __shared__ __align__(128) float win_raw[];
  float* win = align128(win_raw);
  const Tile t = make_tile<C>(tilesX, tilesY);
  const int y = t.ty0 + threadIdx.x / (TW / PXT);
  const int x = t.tx0 + (threadIdx.x % (TW / PXT)) * PXT;
  const bool live = (y < H) && (x < W);
  const size_t pix = ((size_t)t.n * H + (live ? y : 0)) * W + (live ? x : 0);

  float2 f[PXT];
  float gv[PXT * C];
  if (live) {
    const float4 f01 = __ldg(reinterpret_cast<const float4*>(flow + pix * 2));
    const float4 f23 = __ldg(reinterpret_cast<const float4*>(flow + pix * 2) + 1);
    f[0] = make_float2(f01.x, f01.y); f[1] = make_float2(f01.z, f01.w);
    f[2] = make_float2(f23.x, f23.y); f[3] = make_float2(f23.z, f23.w);
#pragma unroll
    for (int v = 0; v < C; ++v) {
      const float4 q = __ldg(reinterpret_cast<const float4*>(gout + pix * C) + v);
      gv[4 * v] = q.x; gv[4 * v + 1] = q.y; gv[4 * v + 2] = q.z; gv[4 * v + 3] = q.w;
    }
  }
  const int sy0 = t.ty0 - Scat<C>::MY, sx0 = t.tx0 - Scat<C>::MX;
  scatter_zero<C>(win);
  __syncthreads();

  if (live) {
    float* gimg = gimage + (size_t)t.n * H * W * C;
#pragma unroll
    for (int i = 0; i < PXT; ++i) {
      const Taps tp = make_taps(f[i].x, f[i].y, x + i, y, W, H);
      if (tp.in) scatter_taps<C>(tp, gv + i * C, win, sy0, sx0, gimg, W);
    }
  }
  scatter_flush<C>(win, gimage, t.n, H, W, sy0, sx0);
}

// ------------------------------------------------------------------------------------- backward: both gradients
// One kernel for the whole FlowWarpGrad op: flow and gradients are read ONCE and kept in registers; the shared-memory
// window first holds the image neighbourhood (TMA) for the flow-gradient gather, then -- zeroed -- collects the
// image-gradient scatter and is flushed with bulk reduce-adds.  gimage must already be zero.
template <int C>
__global__ void __launch_bounds__(NT, 2) warp_tile_bwd(const __grid_constant__ CUtensorMap imap,
                                                       const float* __restrict__ image,
                                                       const float* __restrict__ flow,
                                                       const float* __restrict__ gout, float* __restrict__ gflow,
                                                       float* __restrict__ gimage, int H, int W, int tilesX,
                                                       int tilesY) {
  extern __shared__ __align__(128) float win_raw[];
  __shared__ __align__(8) uint64_t bar;
  float* win = align128(win_raw);
  const Tile t = make_tile<C>(tilesX, tilesY);
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    mbar_expect_tx(smem_u32(&bar), Win<C>::TX_BYTES);
#pragma unroll
    for (int b = 0; b < Win<C>::NB; ++b)
      tma_load_3d(smem_u32(win + b * Win<C>::BOX_FLOATS), &imap, smem_u32(&bar), t.wx0 * C + b * Win<C>::BW, t.wy0,
                  t.n);
  }
  __syncthreads();
  const int y = t.ty0 + threadIdx.x / (TW / PXT);
  const int x = t.tx0 + (threadIdx.x % (TW / PXT)) * PXT;
  const bool live = (y < H) && (x < W);
  const size_t pix = ((size_t)t.n * H + (live ? y : 0)) * W + (live ? x : 0);

  float2 f[PXT];
  float gv[PXT * C];
#pragma unroll
  for (int i = 0; i < PXT; ++i) f[i] = make_float2(0.f, 0.f);
#pragma unroll
  for (int i = 0; i < PXT * C; ++i) gv[i] = 0.f;
  if (live) {
    const float4 f01 = __ldg(reinterpret_cast<const float4*>(flow + pix * 2));
    const float4 f23 = __ldg(reinterpret_cast<const float4*>(flow + pix * 2) + 1);
    f[0] = make_float2(f01.x, f01.y); f[1] = make_float2(f01.z, f01.w);
    f[2] = make_float2(f23.x, f23.y); f[3] = make_float2(f23.z, f23.w);
#pragma unroll
    for (int v = 0; v < C; ++v) {
      const float4 q = __ldg(reinterpret_cast<const float4*>(gout + pix * C) + v);
      gv[4 * v] = q.x; gv[4 * v + 1] = q.y; gv[4 * v + 2] = q.z; gv[4 * v + 3] = q.w;
    }
  }
  mbar_wait(smem_u32(&bar), 0);

  // ---- phase 1: flow gradient (flow_warp_grad.cu.cc:55-86), taps from the staged image window ----
  if (live) {
    const float* img = image + (size_t)t.n * H * W * C;
    float2 gf[PXT];
#pragma unroll
    for (int i = 0; i < PXT; ++i) {
      const Taps tp = make_taps(f[i].x, f[i].y, x + i, y, W, H);
      float bx = 0.f, by = 0.f;
      if (tp.in) {
        float iTL[C], iTR[C], iBL[C], iBR[C];
        const bool inwin = (tp.T >= t.wy0) && (tp.B < t.wy0 + Win<C>::ROWS) && (tp.L >= t.wx0) &&
                           (tp.R < t.wx0 + Win<C>::COLS);
        if (inwin) {
          const float* pL = win + Win<C>::at(tp.T - t.wy0, tp.L - t.wx0);
          const float* pR = win + Win<C>::at(tp.T - t.wy0, tp.R - t.wx0);
          const int dB = (tp.B - tp.T) * Win<C>::BW;
#pragma unroll
          for (int c = 0; c < C; ++c) { iTL[c] = pL[c]; iTR[c] = pR[c]; iBL[c] = pL[dB + c]; iBR[c] = pR[dB + c]; }
        } else {
          const float* pT = img + ((size_t)tp.T * W + tp.L) * C;
          const size_t dR = (size_t)(tp.R - tp.L) * C, dB = (size_t)(tp.B - tp.T) * W * C;
#pragma unroll
          for (int c = 0; c < C; ++c) {
            iTL[c] = __ldg(pT + c); iTR[c] = __ldg(pT + dR + c);
            iBL[c] = __ldg(pT + dB + c); iBR[c] = __ldg(pT + dB + dR + c);
          }
        }
        const float gy = float(tp.B) - tp.y2;  // flow_warp_grad.cu.cc:55
        const float gx = float(tp.R) - tp.x2;  // :72
#pragma unroll
        for (int c = 0; c < C; ++c) {
          float tx = gy * (iTR[c] - iTL[c]);
          tx += (1 - gy) * (iBR[c] - iBL[c]);
          bx += gv[i * C + c] * tx;
          float ty = gx * (iBL[c] - iTL[c]);
          ty += (1 - gx) * (iBR[c] - iTR[c]);
          by += gv[i * C + c] * ty;
        }
      }
      gf[i] = make_float2(bx, by);
    }
    reinterpret_cast<float4*>(gflow + pix * 2)[0] = make_float4(gf[0].x, gf[0].y, gf[1].x, gf[1].y);
    reinterpret_cast<float4*>(gflow + pix * 2)[1] = make_float4(gf[2].x, gf[2].y, gf[3].x, gf[3].y);
  }
  __syncthreads();   // everyone has read the image window

  // ---- phase 2: image gradient (flow_warp_grad.cu.cc:43-53): a (smaller) window in the same buffer, zeroed, collects
  // the scatter ----
  const int sy0 = t.ty0 - Scat<C>::MY, sx0 = t.tx0 - Scat<C>::MX;
  scatter_zero<C>(win);
  __syncthreads();
  if (live) {
    float* gimg = gimage + (size_t)t.n * H * W * C;
#pragma unroll
    for (int i = 0; i < PXT; ++i) {
      const Taps tp = make_taps(f[i].x, f[i].y, x + i, y, W, H);
      if (tp.in) scatter_taps<C>(tp, gv + i * C, win, sy0, sx0, gimg, W);
    }
  }
  scatter_flush<C>(win, gimage, t.n, H, W, sy0, sx0);
}

template <int C>
int launch_tile(int mode, const float* image, const float* flow, const float* gout, float* out, float* out2, int N,
                int H, int W, cudaStream_t st) {
  static PerDeviceOnce once;   // per template instantiation; function attributes are per device
  const cudaError_t err = per_device_once(once, [] {
    const int bytes = (int)Win<C>::BYTES;
    cudaError_t e = cudaFuncSetAttribute(warp_tile_gather<C, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(warp_tile_gather<C, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(warp_tile_gather<C, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(warp_tile_scatter<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(warp_tile_bwd<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    return e;
  });
  if (err != cudaSuccess) { note_cuda_error(err); (void)cudaGetLastError(); return FN2_E_CUDA; }
  const int tilesX = (W + TW - 1) / TW, tilesY = (H + TH - 1) / TH;
  const long long blocks = (long long)N * tilesX * tilesY;
  if (blocks > 0x7fffffffLL) return FN2_E_BAD_SHAPE;
  const size_t smem = Win<C>::BYTES;
  CUtensorMap map;   // over `image`, for the gathers
  if (mode != 2 && !make_row_map(&map, image, N, H, W, C, Win<C>::BW, Win<C>::ROWS)) return FN2_E_CUDA;
  if (mode == 0)
    warp_tile_gather<C, 0><<<(unsigned)blocks, NT, smem, st>>>(map, image, flow, nullptr, out, nullptr, H, W, tilesX,
                                                               tilesY);
  else if (mode == 1)
    warp_tile_gather<C, 1><<<(unsigned)blocks, NT, smem, st>>>(map, image, flow, gout, out, nullptr, H, W, tilesX,
                                                               tilesY);
  else if (mode == 4)   // out = gflow, out2 = gimage
    warp_tile_bwd<C><<<(unsigned)blocks, NT, smem, st>>>(map, image, flow, gout, out, out2, H, W, tilesX, tilesY);
  else if (mode == 3)
    warp_tile_gather<C, 2><<<(unsigned)blocks, NT, smem, st>>>(map, image, flow, gout, out, out2, H, W, tilesX, tilesY);
  else
    warp_tile_scatter<C><<<(unsigned)blocks, NT, smem, st>>>(flow, gout, out, H, W, tilesX, tilesY);
  return after_launch();
}

int dispatch_tile(int mode, const float* image, const float* flow, const float* gout, float* out, int N, int H, int W,
                  int C, cudaStream_t st, float* out2 = nullptr) {
  switch (C) {
    case 1: return launch_tile<1>(mode, image, flow, gout, out, out2, N, H, W, st);
    case 2: return launch_tile<2>(mode, image, flow, gout, out, out2, N, H, W, st);
    case 3: return launch_tile<3>(mode, image, flow, gout, out, out2, N, H, W, st);
    case 4: return launch_tile<4>(mode, image, flow, gout, out, out2, N, H, W, st);
    default: return FN2_E_BAD_SHAPE;
  }
}

}  // namespace

bool flow_warp_tile_supported(int H, int W, int C) { return C >= 1 && C <= 4 && (W % 4) == 0 && H >= 1; }

int flow_warp_tile_forward(const float* image, const float* flow, float* out, int N, int H, int W, int C,
                           cudaStream_t st) {
  return dispatch_tile(0, image, flow, nullptr, out, N, H, W, C, st);
}

// warp + brightness error in one pass (mode 3 of launch_tile = MODE 2 of the gather)
int flow_warp_tile_forward_error(const float* image, const float* flow, const float* reference, float* out, float* err,
                                 int N, int H, int W, int C, cudaStream_t st) {
  return dispatch_tile(3, image, flow, reference, out, N, H, W, C, st, err);
}

// gimage must already be zero (flow_warp_backward memsets it on the same stream)
int flow_warp_tile_backward(const float* image, const float* flow, const float* gout, float* gimage, float* gflow,
                            int N, int H, int W, int C, cudaStream_t st) {
  if (!env_switches().warp_bwd_split) return dispatch_tile(4, image, flow, gout, gflow, N, H, W, C, st, gimage);
  const int rc = dispatch_tile(1, image, flow, gout, gflow, N, H, W, C, st);
  if (rc != FN2_OK) return rc;
  return dispatch_tile(2, nullptr, flow, gout, gimage, N, H, W, C, st);
}

}  // namespace fn2
