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

__device__ __forceinline__ float nan_f() { return __int_as_float(0x7fffffff); }

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


// ---- fast path ------------------------------------------------------------------------------------------------
// Output column nearest to input column x (same float expression wherever it is needed, so it is consistent).
__device__ __forceinline__ int nearest_out_col(int x, float ws) { return (int)rintf((float)x / ws); }

template <int CT>
__global__ void __launch_bounds__(CT <= 2 ? 1024 : 512) downsample_strip_kernel(DsParams p, const float* __restrict__ in,
                                                                float* __restrict__ out) {
  constexpr int NV = 3 * CT;              // sums per slot: value, weight, nan for each channel
  extern __shared__ float sm[];           // [3 slots * NV][blockDim.x] column sums, then [oW][NV] row-chunk partials
  cg::cluster_group cluster = cg::this_cluster();
  const int rs = (int)cluster.num_blocks();
  const int rank = (int)cluster.block_rank();
  const int dy = blockIdx.y, n = blockIdx.z;
  const int x = threadIdx.x, nthr = blockDim.x;
  const float srcy = ((float)dy / (float)(p.oH - 1)) * (float)(p.iH - 1);
  const int isy = (int)roundf(srcy);
  const int ylo = max(isy - p.hr, 0), yhi = min(isy + p.hr, p.iH - 1);
  const int chunk = (yhi - ylo + 1 + rs - 1) / rs;
  const int yb = ylo + rank * chunk, ye = min(yhi, yb + chunk - 1);

  // ---- per-column setup: three candidate output columns and their x-weights ----
  float wx[3];
  const int dxc = nearest_out_col(x, p.ws);
#pragma unroll
  for (int s = 0; s < 3; ++s) {
    const int dx = dxc - 1 + s;
    wx[s] = 0.f;
    if (x < p.iW && dx >= 0 && dx < p.oW) {
      const float srcx = ((float)dx / (float)(p.oW - 1)) * (float)(p.iW - 1);
      const int isx = (int)roundf(srcx);
      if (x >= isx - p.wr && x <= isx + p.wr) wx[s] = fmaxf(0.0f, 1.0f - (fabsf((float)x - srcx) / p.ws));
    }
  }
  float av[3][CT], aw[3][CT], an[3][CT];
#pragma unroll
  for (int s = 0; s < 3; ++s)
#pragma unroll
    for (int c = 0; c < CT; ++c) av[s][c] = aw[s][c] = an[s][c] = 0.f;

  // ---- walk the rows of this CTA's chunk, 4 rows in flight ----
  const float* col = in + ((size_t)n * p.iH * p.iW + (size_t)min(x, p.iW - 1)) * CT;
  constexpr int U = 4;
  for (int y0 = yb; y0 <= ye; y0 += U) {
    float v[U][CT];
    float wy[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int y = min(y0 + u, ye);
      const float* ps = col + (size_t)y * p.iW * CT;
      if constexpr (CT == 2) {
        const float2 t = __ldg(reinterpret_cast<const float2*>(ps));
        v[u][0] = t.x; v[u][1] = t.y;
      } else if constexpr (CT == 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(ps));
        v[u][0] = t.x; v[u][1] = t.y; v[u][2] = t.z; v[u][3] = t.w;
      } else {
#pragma unroll
        for (int c = 0; c < CT; ++c) v[u][c] = __ldg(ps + c);
      }
      wy[u] = (y0 + u <= ye) ? fmaxf(0.0f, 1.0f - (fabsf((float)y - srcy) / p.hs)) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const bool live = (y0 + u <= ye);
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const float w = wx[s] * wy[u];
#pragma unroll
        for (int c = 0; c < CT; ++c) {
          const float sv = v[u][c];
          const bool isnan_ = (sv != sv);
          if (live) {
            an[s][c] += isnan_ ? w : 0.f;
            av[s][c] = fmaf(isnan_ ? 0.f : sv, isnan_ ? 0.f : w, av[s][c]);
            aw[s][c] += isnan_ ? 0.f : w;
          }
        }
      }
    }
  }

  // ---- column sums -> shared, then one warp per output pixel adds its columns in a fixed order ----
#pragma unroll
  for (int s = 0; s < 3; ++s)
#pragma unroll
    for (int c = 0; c < CT; ++c) {
      sm[((s * NV) + c) * nthr + x] = av[s][c];
      sm[((s * NV) + CT + c) * nthr + x] = aw[s][c];
      sm[((s * NV) + 2 * CT + c) * nthr + x] = an[s][c];
    }
  __syncthreads();
  float* part = sm + (size_t)3 * NV * nthr;   // [oW][NV]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = nthr >> 5;
  for (int dx = warp; dx < p.oW; dx += nwarp) {
    // columns whose nearest output column is dx-1, dx or dx+1 (a superset, filtered below)
    const int xlo = max(0, (int)floorf(((float)dx - 1.5f) * p.ws) - 1);
    const int xhi = min(p.iW - 1, (int)ceilf(((float)dx + 1.5f) * p.ws) + 1);
    float tot[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) tot[k] = 0.f;
    for (int xx = xlo + lane; xx <= xhi; xx += 32) {
      const int s = dx - nearest_out_col(xx, p.ws) + 1;
      if (s >= 0 && s < 3) {
#pragma unroll
        for (int k = 0; k < NV; ++k) tot[k] += sm[(s * NV + k) * nthr + xx];
      }
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1)
#pragma unroll
      for (int k = 0; k < NV; ++k) tot[k] += __shfl_xor_sync(FULL, tot[k], off);
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < NV; ++k) part[dx * NV + k] = tot[k];
    }
  }
  cluster.sync();   // every rank's partials are in its shared memory (also a CTA barrier)
  if (rank == 0) {
    for (int idx = threadIdx.x; idx < p.oW * CT; idx += nthr) {
      const int dx = idx / CT, c = idx - dx * CT;
      float v_ = 0.f, w_ = 0.f, n_ = 0.f;
      for (int r = 0; r < rs; ++r) {
        const float* rp = cluster.map_shared_rank(part, r) + dx * NV;
        v_ += rp[c]; w_ += rp[CT + c]; n_ += rp[2 * CT + c];
      }
      out[(((size_t)n * p.oH + dy) * p.oW + dx) * CT + c] = (n_ / w_ > 0.5f) ? nan_f() : v_ / w_;
    }
  }
  cluster.sync();   // keep remote shared memory alive until rank 0 has read it
}

template <int CT>
int launch_strip(const DsParams& p, const float* in, float* out, cudaStream_t st) {
  const int nthr = ((p.iW + 31) / 32) * 32;
  const int rows = 2 * p.hr + 1;
  int rs = 1;
  while (rs < 8 && rows / (rs * 2) >= 8) rs *= 2;   // >= 8 rows per CTA, at most 8 CTAs per cluster
  const size_t smem = ((size_t)9 * CT * nthr + (size_t)p.oW * 3 * CT) * sizeof(float);
  auto kern = downsample_strip_kernel<CT>;
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [&] {
    attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  });
  if (attr_err != cudaSuccess) { note_cuda_error(attr_err); return FN2_E_CUDA; }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)rs, (unsigned)p.oH, (unsigned)p.N);
  cfg.blockDim = dim3((unsigned)nthr);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)rs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p, in, out);
  if (e != cudaSuccess) { note_cuda_error(e); (void)cudaGetLastError(); return FN2_E_CUDA; }
  return after_launch();
}

bool strip_supported(const DsParams& p) {
  if (p.C > MAXC || p.iW > (p.C <= 2 ? 1024 : 512) || p.oH > 65535 || p.N > 65535) return false;
  const size_t smem = ((size_t)9 * p.C * (((p.iW + 31) / 32) * 32) + (size_t)p.oW * 3 * p.C) * sizeof(float);
  return smem <= 200 * 1024;
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
  if (strip_supported(p)) {
    switch (C) {
      case 1: return launch_strip<1>(p, in, out, st);
      case 2: return launch_strip<2>(p, in, out, st);
      case 3: return launch_strip<3>(p, in, out, st);
      default: return launch_strip<4>(p, in, out, st);
    }
  }
  const long long taps = (long long)(2 * p.wr + 1) * (2 * p.hr + 1);
  if (taps <= 64) return launch_group<8>(p, in, out, st);
  if (taps <= 2048) return launch_group<32>(p, in, out, st);
  return launch_group<256>(p, in, out, st);
}

}  // namespace fn2
