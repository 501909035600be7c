// downsample.cu -- NaN-aware triangular-weighted area downsample (forward only, as in the reference).
//
// Semantics follow downsample_kernel_gpu.cu.cc:35-76,92-96: per output (n,dy,dx,c)
//   srcx = (dx/(oW-1))*(iW-1), isrcx = round-half-away(srcx); window [isrc-r, isrc+r] clipped to the image,
//   w = max(0, 1-|xoff-srcx|/ws) * max(0, 1-|yoff-srcy|/hs); NaN samples move their weight to accum_nan;
//   result NaN (0x7fffffff) if accum_nan/accum_weight > 0.5 else accum_value/accum_weight.
// The reference gives each output element to ONE thread that walks its whole window serially (22 785 taps at
// 6x8 from 384x512).  Here a group of G threads (a sub-warp, a warp or a whole CTA, picked from the window size)
// shares one output pixel: the taps are dealt to the group row by row so that consecutive lanes read consecutive
// input pixels (coalesced, all C channels of a pixel by one lane), and the three sums are reduced with shuffles.
// HBM-bound: every input byte is read once from DRAM (re-reads of overlapping windows hit L2).
//
// Fast path (C <= 4, in_w <= 1024): downsample_strip_kernel.  The weight is separable, w = wx(x) * wy(y), and a tap
// with w == 0 adds exact zeros to all three sums, so only the <= 2 (3 with rounding slack) output columns whose
// triangle covers an input column matter.  One CTA owns one output row (n, dy): thread x walks input column x down
// the rows of the window with fully coalesced row reads (in_w * C floats per row), keeping value/weight/NaN sums
// for its three candidate output columns in registers; a deterministic shared-memory pass then sums the columns of
// each output pixel.  Coarse scales (windows up to 155 rows) split the rows over a thread-block cluster of up to 8
// CTAs and combine the partial sums through distributed shared memory, so every scale launches >= ~1500 CTAs at
// batch 32 and each input row is read twice (from L2), not (2*hr+1)*(2*wr+1)/(hs*ws) times with 2 divisions per tap.
#include <cooperative_groups.h>
#include <mutex>
#include "fn2_common.cuh"
#include "fn2_downsample.cuh"

namespace fn2 {
namespace {

namespace cg = cooperative_groups;

constexpr unsigned FULL = 0xffffffffu;
constexpr int MAXC = 4;

struct DsParams {
  int N, iH, iW, C, oH, oW;
  float ws, hs;
  int wr, hr;
};


// G threads per output pixel, G in {8, 32, 256}; CTA = 256 threads.  C <= MAXC channels handled per lane.
template <int G, int CT>
__global__ void __launch_bounds__(256) downsample_kernel(DsParams p, const float* __restrict__ in,
                                                         float* __restrict__ out) {
  __shared__ float red[3 * MAXC][8];
  const int sub = threadIdx.x % G;
  const long long opix = (long long)blockIdx.x * (256 / G) + threadIdx.x / G;
  const long long nopix = (long long)p.N * p.oH * p.oW;
  const bool live = opix < nopix;
  float av[CT], aw[CT], an[CT];
#pragma unroll
  for (int c = 0; c < CT; ++c) av[c] = aw[c] = an[c] = 0.f;

  int dx = 0, dy = 0;
  long long n = 0;
  if (live) {
    dx = (int)(opix % p.oW);
    dy = (int)((opix / p.oW) % p.oH);
    n = opix / ((long long)p.oW * p.oH);
    const float srcx = ((float)dx / (float)(p.oW - 1)) * (float)(p.iW - 1);
    const float srcy = ((float)dy / (float)(p.oH - 1)) * (float)(p.iH - 1);
    const int isx = (int)roundf(srcx), isy = (int)roundf(srcy);
    const int x0 = max(isx - p.wr, 0), x1 = min(isx + p.wr, p.iW - 1);
    const int y0 = max(isy - p.hr, 0), y1 = min(isy + p.hr, p.iH - 1);
    const int ww = x1 - x0 + 1;
    const int taps = ww * (y1 - y0 + 1);
    const float* src = in + (size_t)n * p.iH * p.iW * p.C;
    for (int t = sub; t < taps; t += G) {
      const int yoff = y0 + t / ww, xoff = x0 + t % ww;
      const float w = fmaxf(0.0f, 1.0f - (fabsf((float)xoff - srcx) / p.ws)) *
                      fmaxf(0.0f, 1.0f - (fabsf((float)yoff - srcy) / p.hs));
      const float* ps = src + ((size_t)yoff * p.iW + xoff) * p.C;
#pragma unroll
      for (int c = 0; c < CT; ++c) {
        if (c < p.C) {
          float s = __ldg(ps + c);
          float wc = w;
          if (s != s) { an[c] += w; s = 0.f; wc = 0.f; }
          av[c] = fmaf(s, wc, av[c]);
          aw[c] += wc;
        }
      }
    }
  }
  // reduce across the group
  constexpr int WG = G < 32 ? G : 32;
#pragma unroll
  for (int off = WG / 2; off >= 1; off >>= 1) {
#pragma unroll
    for (int c = 0; c < CT; ++c) {
      av[c] += __shfl_xor_sync(FULL, av[c], off);
      aw[c] += __shfl_xor_sync(FULL, aw[c], off);
      an[c] += __shfl_xor_sync(FULL, an[c], off);
    }
  }
  if (G == 256) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
#pragma unroll
      for (int c = 0; c < CT; ++c) { red[c][warp] = av[c]; red[MAXC + c][warp] = aw[c]; red[2 * MAXC + c][warp] = an[c]; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
      for (int c = 0; c < CT; ++c) {
        av[c] = aw[c] = an[c] = 0.f;
        for (int w8 = 0; w8 < 8; ++w8) { av[c] += red[c][w8]; aw[c] += red[MAXC + c][w8]; an[c] += red[2 * MAXC + c][w8]; }
      }
    }
  }
  if (live && sub == 0) {
    float* po = out + (size_t)opix * p.C;
#pragma unroll
    for (int c = 0; c < CT; ++c)
      if (c < p.C) po[c] = (an[c] / aw[c] > 0.5f) ? nan_f() : av[c] / aw[c];
  }
}

// Any channel count: one thread per output element, channel innermost (coalesced), window walked serially.
__global__ void __launch_bounds__(256) downsample_wide_kernel(DsParams p, const float* __restrict__ in,
                                                              float* __restrict__ out) {
  const long long total = (long long)p.N * p.oH * p.oW * p.C;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(idx % p.C);
    const int dx = (int)((idx / p.C) % p.oW);
    const int dy = (int)((idx / p.C / p.oW) % p.oH);
    const long long n = idx / p.C / p.oW / p.oH;
    const float srcx = ((float)dx / (float)(p.oW - 1)) * (float)(p.iW - 1);
    const float srcy = ((float)dy / (float)(p.oH - 1)) * (float)(p.iH - 1);
    const int isx = (int)roundf(srcx), isy = (int)roundf(srcy);
    float av = 0.f, aw = 0.f, an = 0.f;
    for (int yoff = max(isy - p.hr, 0); yoff <= min(isy + p.hr, p.iH - 1); ++yoff)
      for (int xoff = max(isx - p.wr, 0); xoff <= min(isx + p.wr, p.iW - 1); ++xoff) {
        float s = __ldg(in + (((size_t)n * p.iH + yoff) * p.iW + xoff) * p.C + c);
        float w = fmaxf(0.0f, 1.0f - (fabsf((float)xoff - srcx) / p.ws)) *
                  fmaxf(0.0f, 1.0f - (fabsf((float)yoff - srcy) / p.hs));
        if (s != s) { an += w; s = 0.f; w = 0.f; }
        av = fmaf(s, w, av);
        aw += w;
      }
    out[idx] = (an / aw > 0.5f) ? nan_f() : av / aw;
  }
}


// ---- fast path: row-band kernel ---------------------------------------------------------------------------------
// The weight is separable, w = wx(x) * wy(y), and a tap with w == 0 adds exact zeros to all three sums.  The triangle
// of output row dy is non-zero only for |y - srcy(dy)| < hs and the centres srcy(dy) are hs apart, so an input row
// feeds at most TWO output rows (dlo = the last dy with srcy(dy) <= y, and dlo + 1); same along x.  The window clip of
// the reference ([isrc - r, isrc + r], r = ceil(scale)) never removes a tap of non-zero weight.
//
// One CTA (or a cluster of rs CTAs that split the rows) owns a group of G consecutive output rows of one image.
// Thread x owns input column x: it walks down the rows of the band with coalesced row reads, U rows in flight,
// keeping the vertical sums (value, weight, NaN weight per channel) of the two live output rows in registers.  When
// the walk passes srcy(dlo + 1) the finished row is flushed: every thread writes its column sums, pre-multiplied by
// its two x-weights, to shared memory, and the (dx, sum) pairs are then added up by sub-warp groups in a fixed order.
// Cluster ranks combine their partial rows through distributed shared memory.  Every input byte is read once from
// DRAM (the (G+1)/G re-read of shared band rows hits L2), and the per-tap work is 3 FMAs per channel and live row.
struct BandPlan {
  int G, rs, px;
};

constexpr int DS_MAX_CHUNK = 256;   // rows per CTA the row table can hold
constexpr int DS_MAX_OW = 512;      // output columns the range table can hold
constexpr int DS_SMS = 148;         // B200

// CT channels, PX adjacent pixels per thread (CT * PX floats per row load, 128-bit when that is a multiple of 4),
// U rows in flight per thread.
template <int CT, int PX, int U>
__global__ void __launch_bounds__(512, 2) downsample_band_kernel(DsParams p, int G, const float* __restrict__ in,
                                                                 float* __restrict__ out) {
  constexpr int NV = 2 * CT;                  // per column and output row: value and weight sums per channel
  constexpr int NF = CT * PX;                 // floats per thread and row
  extern __shared__ float sm[];
  __shared__ float rowA[DS_MAX_CHUNK], rowB[DS_MAX_CHUNK];
  __shared__ int rowD[DS_MAX_CHUNK];
  __shared__ int colStart[DS_MAX_OW + 1];     // first input column whose nearest-left output column is dx
  cg::cluster_group cluster = cg::this_cluster();
  const int rs = (int)cluster.num_blocks();
  const int rank = (int)cluster.block_rank();
  const int n = blockIdx.z;
  const int d0 = blockIdx.y * G, d1 = min(p.oH, d0 + G);
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int ncols = nthr * PX;
  float* sV = sm;                                    // [G][NV][ncols] vertical sums of the finished output rows
  float* sTot = sV + (size_t)G * NV * ncols;         // [G] total vertical weight of each output row (NaN or not)
  float* wA = sTot + ((G + 3) & ~3);                 // [ncols] x-weight towards output column dxl(x)
  float* wB = wA + ncols;                            // [ncols] x-weight towards dxl(x) + 1
  float* part = wB + ncols;                          // [G][oW][NV + 1] partial output rows (clusters only)

  // ---- rows of the band, and this rank's chunk of them ----
  const float sy0 = src_pos(d0, p.oH - 1, p.iH - 1), sy1 = src_pos(d1 - 1, p.oH - 1, p.iH - 1);
  const int ylo = max(0, (int)floorf(sy0 - p.hs)), yhi = min(p.iH - 1, (int)ceilf(sy1 + p.hs));
  const int chunk = (yhi - ylo + 1 + rs - 1) / rs;
  const int yb = ylo + rank * chunk, ye = min(yhi, yb + chunk - 1);

  if (rs > 1) {   // a rank only finishes the rows its chunk reaches: the others must read as zero
    for (int i = tid; i < G * NV * ncols; i += nthr) sV[i] = 0.f;
    for (int i = tid; i < G; i += nthr) sTot[i] = 0.f;
  }
  for (int i = tid; i <= ye - yb; i += nthr) {
    const int y = yb + i;
    const int dlo = last_centre_le(y, p.hs, p.oH, p.iH);
    const bool okA = dlo >= d0 && dlo < d1, okB = dlo + 1 >= d0 && dlo + 1 < d1;
    rowA[i] = okA ? tri((float)y, src_pos(dlo, p.oH - 1, p.iH - 1), p.hs) : 0.f;
    rowB[i] = okB ? tri((float)y, src_pos(dlo + 1, p.oH - 1, p.iH - 1), p.hs) : 0.f;
    rowD[i] = dlo;
  }
  for (int dx = tid; dx <= p.oW; dx += nthr)
    colStart[dx] = dx == 0 ? 0 : (dx == p.oW ? p.iW : (int)ceilf(src_pos(dx, p.oW - 1, p.iW - 1)));
  const int x0 = tid * PX;
#pragma unroll
  for (int q = 0; q < PX; ++q) {
    const int x = x0 + q;
    const bool colok = x < p.iW;
    const int dxl = last_centre_le(min(x, p.iW - 1), p.ws, p.oW, p.iW);
    wA[x] = colok ? tri((float)x, src_pos(dxl, p.oW - 1, p.iW - 1), p.ws) : 0.f;
    wB[x] = (colok && dxl + 1 < p.oW) ? tri((float)x, src_pos(dxl + 1, p.oW - 1, p.iW - 1), p.ws) : 0.f;
  }
  __syncthreads();

  float av[2][NF], aw[2][NF], tot[2];
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    tot[s] = 0.f;
#pragma unroll
    for (int f = 0; f < NF; ++f) av[s][f] = aw[s][f] = 0.f;
  }
  // a finished output row: park the column sums of slot 0 in shared memory (each thread its own columns)
  auto flush = [&](int dy) {
    if (dy < d0 || dy >= d1) return;
    float* row = sV + (size_t)(dy - d0) * NV * ncols;
#pragma unroll
    for (int q = 0; q < PX; ++q)
#pragma unroll
      for (int c = 0; c < CT; ++c) {
        row[c * ncols + x0 + q] = av[0][q * CT + c];
        row[(CT + c) * ncols + x0 + q] = aw[0][q * CT + c];
      }
    if (tid == 0) sTot[dy - d0] = tot[0];
  };
  auto shift = [&]() {
    tot[0] = tot[1]; tot[1] = 0.f;
#pragma unroll
    for (int f = 0; f < NF; ++f) { av[0][f] = av[1][f]; aw[0][f] = aw[1][f]; av[1][f] = aw[1][f] = 0.f; }
  };

  // ---- walk the rows ----
  const float* col = in + ((size_t)n * p.iH * p.iW + (size_t)min(x0, p.iW - PX)) * CT;
  int cur = (yb <= ye) ? rowD[0] : 0;
  for (int y0 = yb; y0 <= ye; y0 += U) {
    float v[U][NF];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const float* ps = col + (size_t)min(y0 + u, ye) * p.iW * CT;
      if constexpr (NF % 4 == 0) {
#pragma unroll
        for (int f = 0; f < NF; f += 4) {
          const float4 t = __ldg(reinterpret_cast<const float4*>(ps + f));
          v[u][f] = t.x; v[u][f + 1] = t.y; v[u][f + 2] = t.z; v[u][f + 3] = t.w;
        }
      } else if constexpr (NF % 2 == 0) {
#pragma unroll
        for (int f = 0; f < NF; f += 2) {
          const float2 t = __ldg(reinterpret_cast<const float2*>(ps + f));
          v[u][f] = t.x; v[u][f + 1] = t.y;
        }
      } else {
#pragma unroll
        for (int f = 0; f < NF; ++f) v[u][f] = __ldg(ps + f);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (y0 + u <= ye) {                                  // uniform
        const int i = y0 + u - yb;
        const int dlo = rowD[i];
        if (dlo != cur) {                                  // uniform: the walk passed srcy(cur + 1)
          flush(cur);
          shift();
          cur = dlo;
        }
        const float wa = rowA[i], wb = rowB[i];
        tot[0] += wa; tot[1] += wb;
#pragma unroll
        for (int f = 0; f < NF; ++f) {
          const float sv = v[u][f];
          if (sv == sv) {                                  // NaN samples only count in the total weight
            av[0][f] = fmaf(sv, wa, av[0][f]); aw[0][f] += wa;
            av[1][f] = fmaf(sv, wb, av[1][f]); aw[1][f] += wb;
          }
        }
      }
    }
  }
  if (yb <= ye) {
    flush(cur);
    shift();
    flush(cur + 1);
  }
  __syncthreads();

  // ---- horizontal pass: J threads per output pixel (dyl, dx) add its columns in a fixed order ----
  const int items = (d1 - d0) * p.oW;
  int J = 1;
  while (J < 32 && J * 2 * items <= nthr) J *= 2;
  const int j = tid % J;
  for (int it0 = 0; it0 < items; it0 += nthr / J) {
    const int item = it0 + tid / J;
    const bool live = item < items;
    const int dyl = live ? item / p.oW : 0, dx = live ? item - dyl * p.oW : 0;
    float acc[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) acc[k] = 0.f;
    float wsum = 0.f;                                        // sum of the x-weights (for the total weight)
    if (live) {
      const float* row = sV + (size_t)dyl * NV * ncols;
      const int c1 = colStart[dx], c2 = colStart[dx + 1];
      if (dx > 0)
        for (int xx = colStart[dx - 1] + j; xx < c1; xx += J) {   // columns whose right neighbour is dx
          const float w = wB[xx];
          wsum += w;
#pragma unroll
          for (int k = 0; k < NV; ++k) acc[k] = fmaf(w, row[k * ncols + xx], acc[k]);
        }
      for (int xx = c1 + j; xx < c2; xx += J) {                    // columns whose left neighbour is dx
        const float w = wA[xx];
        wsum += w;
#pragma unroll
        for (int k = 0; k < NV; ++k) acc[k] = fmaf(w, row[k * ncols + xx], acc[k]);
      }
    }
    for (int off = J >> 1; off >= 1; off >>= 1) {
      wsum += __shfl_xor_sync(FULL, wsum, off);
#pragma unroll
      for (int k = 0; k < NV; ++k) acc[k] += __shfl_xor_sync(FULL, acc[k], off);
    }
    if (live && j == 0) {
      const float t_ = wsum * sTot[dyl];
      if (rs == 1) {
#pragma unroll
        for (int c = 0; c < CT; ++c) {
          const float n_ = t_ - acc[CT + c];                 // weight of the NaN samples (accum_nan, :61-65)
          out[(((size_t)n * p.oH + d0 + dyl) * p.oW + dx) * CT + c] =
              (n_ / acc[CT + c] > 0.5f) ? nan_f() : acc[c] / acc[CT + c];
        }
      } else {
        float* pr = part + (size_t)item * (NV + 1);
#pragma unroll
        for (int k = 0; k < NV; ++k) pr[k] = acc[k];
        pr[NV] = t_;
      }
    }
  }
  if (rs == 1) return;

  // ---- clusters: combine the ranks' partial rows in rank order and finish ----
  cluster.sync();
  for (int idx = rank * nthr + tid; idx < items * CT; idx += rs * nthr) {
    const int c = idx % CT, item = idx / CT;
    float v_ = 0.f, w_ = 0.f, t_ = 0.f;
    for (int r = 0; r < rs; ++r) {
      const float* rp = cluster.map_shared_rank(part, r) + (size_t)item * (NV + 1);
      v_ += rp[c]; w_ += rp[CT + c]; t_ += rp[NV];
    }
    const float n_ = t_ - w_;
    const int dyl = item / p.oW, dx = item - dyl * p.oW;
    out[(((size_t)n * p.oH + d0 + dyl) * p.oW + dx) * CT + c] = (n_ / w_ > 0.5f) ? nan_f() : v_ / w_;
  }
  cluster.sync();   // keep this CTA's shared memory alive until every rank has read it
}

int band_threads(const DsParams& p, int px) { return (((p.iW + px - 1) / px + 31) / 32) * 32; }

size_t band_smem(const DsParams& p, int G, int rs, int px) {
  const int NV = 2 * p.C;
  const size_t ncols = (size_t)band_threads(p, px) * px;
  return ((size_t)G * NV * ncols + ((G + 3) & ~3) + 2 * ncols + (rs > 1 ? (size_t)G * p.oW * (NV + 1) : 0)) *
         sizeof(float);
}

bool band_plan(const DsParams& p, BandPlan* bp) {
  if (p.C > MAXC || p.oH > 65535 || p.N > 65535 || p.oW > DS_MAX_OW) return false;
  if (!(p.hs >= 1.0f) || !(p.ws >= 1.0f)) return false;   // upsampling: more than two outputs per input row
  int px = 1;                                               // pixels per thread: 16-byte row loads when they stay aligned
  if (p.C == 1 && p.iW % 4 == 0) px = 4;
  if (p.C == 2 && p.iW % 2 == 0) px = 2;
  const int nthr = band_threads(p, px);
  if (nthr > 512) return false;
  // Pick the rows per CTA: cost = (waves of CTAs over the 148 SMs) x (rows walked per CTA + a fixed overhead).
  double best = 1e30;
  int bG = 0, bRs = 1;
  for (int G = 1; G <= p.oH && G <= 64; ++G) {
    const int rows = (int)ceilf((float)(G + 1) * p.hs) + 2;
    if (G > 1 && rows > 64) break;
    int rs = 1;
    while (rs < 8 && rows > 40 * rs) rs *= 2;
    if ((rows + rs - 1) / rs + 2 > DS_MAX_CHUNK) continue;
    const size_t smem = band_smem(p, G, rs, px) + 8 * 1024;   // + static tables and the per-CTA reserve
    if (smem > 200 * 1024) continue;
    int bps = (int)((227 * 1024) / smem);
    bps = bps > 1024 / nthr ? 1024 / nthr : bps;              // 64 registers per thread: 1024 threads per SM
    if (bps < 1) bps = 1;
    const long long ctas = (long long)p.N * ((p.oH + G - 1) / G) * rs;
    const long long slots = (long long)DS_SMS * bps;
    const double waves = (double)((ctas + slots - 1) / slots);
    const double cost = waves * ((double)rows / rs + 10.0) * (bps < 2 ? 1.5 : 1.0);
    if (cost < best) { best = cost; bG = G; bRs = rs; }
  }
  if (bG == 0) return false;
  bp->G = bG; bp->rs = bRs; bp->px = px;
  return true;
}

template <int CT, int PX, int U>
int launch_band(const DsParams& p, const BandPlan& bp, const float* in, float* out, cudaStream_t st) {
  auto kern = downsample_band_kernel<CT, PX, U>;
  static PerDeviceOnce once;   // one per template instantiation; function attributes are per device
  const cudaError_t attr_err = per_device_once(once, [&] {
    return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  });
  if (attr_err != cudaSuccess) { note_cuda_error(attr_err); (void)cudaGetLastError(); return FN2_E_CUDA; }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)bp.rs, (unsigned)((p.oH + bp.G - 1) / bp.G), (unsigned)p.N);
  cfg.blockDim = dim3((unsigned)band_threads(p, PX));
  cfg.dynamicSmemBytes = band_smem(p, bp.G, bp.rs, PX);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)bp.rs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p, bp.G, in, out);
  if (e != cudaSuccess) { note_cuda_error(e); (void)cudaGetLastError(); return FN2_E_CUDA; }
  return after_launch();
}

template <int G>
int launch_group(const DsParams& p, const float* in, float* out, cudaStream_t st) {
  const long long nopix = (long long)p.N * p.oH * p.oW;
  const long long blocks = (nopix + (256 / G) - 1) / (256 / G);
  if (blocks > 0x7fffffffLL) return FN2_E_BAD_SHAPE;
  if (p.C <= 2)
    downsample_kernel<G, 2><<<(unsigned)blocks, 256, 0, st>>>(p, in, out);
  else
    downsample_kernel<G, MAXC><<<(unsigned)blocks, 256, 0, st>>>(p, in, out);
  return after_launch();
}

}  // namespace

int downsample_forward(const float* in, float* out, int N, int iH, int iW, int C, int oH, int oW, cudaStream_t st) {
  DsParams p;
  p.N = N; p.iH = iH; p.iW = iW; p.C = C; p.oH = oH; p.oW = oW;
  p.ws = (float)(iW - 1) / (float)(oW - 1);  // downsample_kernel_gpu.cu.cc:92-93
  p.hs = (float)(iH - 1) / (float)(oH - 1);
  p.wr = (int)ceilf(p.ws);                   // :95-96
  p.hr = (int)ceilf(p.hs);
  if (C > MAXC) {
    const long long total = (long long)N * oH * oW * C;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    downsample_wide_kernel<<<(unsigned)blocks, 256, 0, st>>>(p, in, out);
    return after_launch();
  }
  BandPlan bp;
  if (band_plan(p, &bp)) {
    switch (C) {
      case 1: return bp.px == 4 ? launch_band<1, 4, 4>(p, bp, in, out, st) : launch_band<1, 1, 8>(p, bp, in, out, st);
      case 2: return bp.px == 2 ? launch_band<2, 2, 4>(p, bp, in, out, st) : launch_band<2, 1, 8>(p, bp, in, out, st);
      case 3: return launch_band<3, 1, 4>(p, bp, in, out, st);
      default: return launch_band<4, 1, 4>(p, bp, in, out, st);
    }
  }
  const long long taps = (long long)(2 * p.wr + 1) * (2 * p.hr + 1);
  if (taps <= 64) return launch_group<8>(p, in, out, st);
  if (taps <= 2048) return launch_group<32>(p, in, out, st);
  return launch_group<256>(p, in, out, st);
}

}  // namespace fn2
