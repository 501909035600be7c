// corr_generic.cu -- fp32 correlation for ANY (kernel_size, max_displacement, stride_1, stride_2, pad).
//
// These kernels cover the whole parameter space of the reference op (correlation_kernel.cu.cc:21-119,
// correlation_grad_kernel.cu.cc:20-189); the FlowNetC call (k=1, s1=1) is served by the tiled kernels in
// corr_tc.cu / corr_tc_x3.cu.  Zero padding is a bounds test on the un-padded tensors: no padded temporaries,
// no memset (the reference's Pad, pad.cu.cc:14-74, is gone).  One launch covers the whole batch.
#include "fn2_common.cuh"

namespace fn2 {

namespace {

constexpr unsigned FULL = 0xffffffffu;

// Sum v[i] over the 32 lanes for 32 values at once; lane l ends up with the total of v[l] in v[0].
// 31 shuffles instead of 32 x 5.
__device__ __forceinline__ float transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      float send = up ? v[i] : v[i + off];
      float keep = up ? v[i + off] : v[i];
      v[i] = keep + __shfl_xor_sync(FULL, send, off);
    }
  }
  return v[0];
}

// One warp per output pixel; 32 displacements per pass.
template <bool VEC4>
__global__ void __launch_bounds__(256) corr_generic_fwd_kernel(CorrGeom g, const float* __restrict__ a,
                                                               const float* __restrict__ b,
                                                               float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const long long pix = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long npix = (long long)g.N * g.oH * g.oW;
  if (pix >= npix) return;
  const int xo = (int)(pix % g.oW);
  const int yo = (int)((pix / g.oW) % g.oH);
  const int n = (int)(pix / ((long long)g.oW * g.oH));
  const int ya = yo * g.s1 + g.off, xa = xo * g.s1 + g.off;
  const float* an = a + (size_t)n * g.H * g.W * g.C;
  const float* bn = b + (size_t)n * g.bH * g.W * g.C;   // b may carry row halos: its row = a-row + bTop
  const float sumelems = (float)(g.k * g.k * g.C);
  float* po = out + (size_t)pix * g.oPitch;

  for (int d0 = 0; d0 < g.D; d0 += 32) {
    float part[32];
#pragma unroll
    for (int dd = 0; dd < 32; ++dd) {
      part[dd] = 0.f;
      const int d = d0 + dd;
      if (d < g.D) {
        const int s2o = (d % g.gw - g.r) * g.s2, s2p = (d / g.gw - g.r) * g.s2;
        for (int j = 0; j < g.k; ++j) {
          const int y1 = ya + j, y2 = y1 + s2p + g.bTop;
          if (y1 < 0 || y1 >= g.H || y2 < 0 || y2 >= g.bH) continue;
          for (int i = 0; i < g.k; ++i) {
            const int x1 = xa + i, x2 = x1 + s2o;
            if (x1 < 0 || x1 >= g.W || x2 < 0 || x2 >= g.W) continue;
            const float* pa = an + ((size_t)y1 * g.W + x1) * g.C;
            const float* pb = bn + ((size_t)y2 * g.W + x2) * g.C;
            if (VEC4) {
              for (int c = lane * 4; c < g.C; c += 128) {
                const float4 av = __ldg(reinterpret_cast<const float4*>(pa + c));
                const float4 bv = __ldg(reinterpret_cast<const float4*>(pb + c));
                part[dd] = fmaf(av.x, bv.x, part[dd]);
                part[dd] = fmaf(av.y, bv.y, part[dd]);
                part[dd] = fmaf(av.z, bv.z, part[dd]);
                part[dd] = fmaf(av.w, bv.w, part[dd]);
              }
            } else {
              for (int c = lane; c < g.C; c += 32) part[dd] = fmaf(__ldg(pa + c), __ldg(pb + c), part[dd]);
            }
          }
        }
      }
    }
    const float tot = transpose_reduce32(part, lane);
    if (d0 + lane < g.D) {
      const float v = tot / sumelems;
      po[d0 + lane] = v > 0.f ? v : v * g.slope;
    }
  }
}

// One CTA per pixel of the (possibly taller, row-halo) b-space, threads over channels.  d/da (when the pixel is also
// a pixel of a) and d/db in one pass, no atomics.
__global__ void __launch_bounds__(256) corr_generic_bwd_kernel(CorrGeom g, const float* __restrict__ gout,
                                                               const float* __restrict__ a,
                                                               const float* __restrict__ b, float* __restrict__ ga,
                                                               float* __restrict__ gb) {
  const long long pix = blockIdx.x;  // n*bH*W + yb*W + xx
  const int xx = (int)(pix % g.W);
  const int ybrow = (int)((pix / g.W) % g.bH);
  const int n = (int)(pix / ((long long)g.W * g.bH));
  const int yy = ybrow - g.bTop;             // the same row in a's coordinates (outside [0,H) on halo rows)
  const bool has_a = (yy >= 0) && (yy < g.H);
  const int x = xx + g.pad, y = yy + g.pad;  // padded coordinates
  const float* an = a + (size_t)n * g.H * g.W * g.C;
  const float* bn = b + (size_t)n * g.bH * g.W * g.C;
  const float* gn = gout + (size_t)n * g.oH * g.oW * g.gPitch;
  const float* mn = g.gMask ? g.gMask + (size_t)n * g.oH * g.oW * g.gPitch : nullptr;
  // gradient of output element (y2, x2, op), times the fused LeakyReLU's derivative when a mask is given
  auto gval = [&](int y2, int x2, int op) {
    const size_t i = ((size_t)y2 * g.oW + x2) * g.gPitch + op;
    const float v = __ldg(gn + i);
    return mn ? (__ldg(mn + i) > 0.f ? v : v * g.slope) : v;
  };
  const float sumelems = (float)(g.k * g.k * g.C);

  // output range covering this element for d/da (same for every displacement)
  int axmin = ceil_div_s(x - 2 * g.kr - g.md, g.s1), aymin = ceil_div_s(y - 2 * g.kr - g.md, g.s1);
  int axmax = floor_div(x - g.md, g.s1), aymax = floor_div(y - g.md, g.s1);
  const bool a_any = has_a && (axmax >= 0) && (aymax >= 0) && (axmin <= g.oW - 1) && (aymin <= g.oH - 1);
  axmin = max(0, axmin); axmax = min(g.oW - 1, axmax);
  aymin = max(0, aymin); aymax = min(g.oH - 1, aymax);

  for (int c = threadIdx.x; c < g.C; c += blockDim.x) {
    float sa = 0.f, sb = 0.f;
    for (int p = -g.r; p <= g.r; ++p) {
      for (int o = -g.r; o <= g.r; ++o) {
        const int op = (p + g.r) * g.gw + (o + g.r);
        const int s2o = g.s2 * o, s2p = g.s2 * p;
        if (a_any) {
          const int yb = y + s2p - g.pad + g.bTop, xb = x + s2o - g.pad;
          if (yb >= 0 && yb < g.bH && xb >= 0 && xb < g.W) {
            const float bv = __ldg(bn + ((size_t)yb * g.W + xb) * g.C + c);
            for (int y2 = aymin; y2 <= aymax; ++y2)
              for (int x2 = axmin; x2 <= axmax; ++x2)
                sa = fmaf(gval(y2, x2, op), bv, sa);
          }
        }
        {
          int xmin = ceil_div_s(x - 2 * g.kr - g.md - s2o, g.s1), ymin = ceil_div_s(y - 2 * g.kr - g.md - s2p, g.s1);
          int xmax = floor_div(x - g.md - s2o, g.s1), ymax = floor_div(y - g.md - s2p, g.s1);
          if ((xmax >= 0) && (ymax >= 0) && (xmin <= g.oW - 1) && (ymin <= g.oH - 1)) {
            xmin = max(0, xmin); xmax = min(g.oW - 1, xmax);
            ymin = max(0, ymin); ymax = min(g.oH - 1, ymax);
            const int yq = y - s2p - g.pad, xq = x - s2o - g.pad;
            if (yq >= 0 && yq < g.H && xq >= 0 && xq < g.W) {
              const float av = __ldg(an + ((size_t)yq * g.W + xq) * g.C + c);
              for (int y2 = ymin; y2 <= ymax; ++y2)
                for (int x2 = xmin; x2 <= xmax; ++x2)
                  sb = fmaf(gval(y2, x2, op), av, sb);
            }
          }
        }
      }
    }
    if (has_a) ga[(((size_t)n * g.H + yy) * g.W + xx) * g.C + c] = sa / sumelems;
    gb[(size_t)pix * g.C + c] = sb / sumelems;
  }
}

}  // namespace

int corr_generic_forward(const CorrGeom& g, const float* a, const float* b, float* out, cudaStream_t st) {
  const long long npix = (long long)g.N * g.oH * g.oW;
  const int warps = 8;
  const long long blocks = (npix + warps - 1) / warps;
  if (blocks > 0x7fffffffLL) return FN2_E_BAD_SHAPE;
  if (g.C % 4 == 0)
    corr_generic_fwd_kernel<true><<<(unsigned)blocks, warps * 32, 0, st>>>(g, a, b, out);
  else
    corr_generic_fwd_kernel<false><<<(unsigned)blocks, warps * 32, 0, st>>>(g, a, b, out);
  return after_launch();
}

int corr_generic_backward(const CorrGeom& g, const float* gout, const float* a, const float* b, float* ga,
                          float* gb, cudaStream_t st) {
  const long long npix = (long long)g.N * g.bH * g.W;
  if (npix > 0x7fffffffLL) return FN2_E_BAD_SHAPE;
  int threads = ((g.C + 31) / 32) * 32;
  if (threads > 256) threads = 256;
  corr_generic_bwd_kernel<<<(unsigned)npix, threads, 0, st>>>(g, gout, a, b, ga, gb);
  return after_launch();
}

}  // namespace fn2
