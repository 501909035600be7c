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
#include "fn2_common.cuh"

namespace fn2 {
namespace {

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
  const long long taps = (long long)(2 * p.wr + 1) * (2 * p.hr + 1);
  if (taps <= 64) return launch_group<8>(p, in, out, st);
  if (taps <= 2048) return launch_group<32>(p, in, out, st);
  return launch_group<256>(p, in, out, st);
}

}  // namespace fn2
