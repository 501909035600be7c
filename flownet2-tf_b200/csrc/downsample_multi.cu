// downsample_multi.cu -- all the loss scales of one tensor in ONE pass over the input.
//
// The reference's loss (flownet_c.py:135-160, flownet_s.py / flownet_sd.py likewise) calls downsample(gt_flow, size)
// five times on the same [N,384,512,2] tensor; every call re-reads the whole input.  Each scale is the NaN-aware
// triangular-weighted average of downsample_kernel_gpu.cu.cc:35-76, and every scale is
// computed DIRECTLY from the input (the scale centres are not nested, so no scale can be derived from another).
//
// Work split: a thread-block cluster of up to 16 CTAs (8 once the batch fills the chip) owns one image; rank r walks the input rows
// [r*chunk, (r+1)*chunk).  Thread x owns PX adjacent input columns and reads each of its rows exactly once (coalesced
// 16-byte loads, U rows in flight).  The weight is separable, w = wx(x) * wy(y): an input row feeds at most two output
// rows of every scale (the last centre at or above it, and the next one), so the thread keeps, per scale, the vertical
// value / weight sums of those two live output rows in registers -- 2 FMAs per input float, scale and live row.  When
// the walk passes the centre of the next output row of a scale (the same row index for every thread: a uniform
// branch), the finished row is parked in shared memory and summed horizontally by sub-warp groups in a fixed order
// into the rank's partial-result table.  After the walk the ranks add their partial rows through distributed shared
// memory in rank order (deterministic) and apply the NaN rule (:72-76).  Every input byte is read once from DRAM;
// all scales together write < 2 % of that.
#include <cstdlib>
#include <cooperative_groups.h>
#include <mutex>
#include "fn2_common.cuh"
#include "fn2_downsample.cuh"

namespace fn2 {
namespace {

namespace cg = cooperative_groups;

constexpr unsigned FULL = 0xffffffffu;
constexpr int MS = 5;          // scales per launch (the reference's losses use five, flownet_c.py:135-160)
constexpr int MAX_RANKS = 16;  // largest cluster (non-portable size, opted into per kernel); 8 is the portable fallback

struct MsParams {
  int N, iH, iW, ns, chunk;
  int oH[MS], oW[MS], maxRows[MS];
  int csOff[MS], trOff[MS], ptOff[MS];   // offsets of each scale's column-start / row-total / partial tables
  int csTotal, trTotal, ptTotal;
  float hs[MS], ws[MS];
  float* out[MS];
};

// shared-memory words (4 bytes each) of one CTA
__host__ __device__ inline size_t ms_smem_words(int chunk, int ncols, int nv, int csTotal, int trTotal, int ptTotal) {
  return (size_t)3 * MS * chunk + chunk      // rowA, rowB, rowD, flush mask
         + (size_t)2 * MS * ncols            // wA, wB
         + csTotal + trTotal                 // colStart, totRow
         + (size_t)nv * ncols                // parked row
         + ptTotal;                          // partial results
}

#ifdef FN2_TIMELINE
__device__ long long g_tld[16];
#define TLD(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) g_tld[i] = clock64(); } while (0)
#else
#define TLD(i) do {} while (0)
#endif

template <int CT, int PX, int U>
__global__ void __launch_bounds__(512) downsample_multi_kernel(MsParams p, const float* __restrict__ in) {
  constexpr int NV = 2 * CT;                  // per column and output row: value and weight sums per channel
  constexpr int NF = CT * PX;                 // floats per thread and row
  extern __shared__ float sm[];
  __shared__ int dF[MAX_RANKS][MS], dL[MAX_RANKS][MS];   // output rows [dF, dL] of scale s that rank r feeds
  cg::cluster_group cluster = cg::this_cluster();
  const int rs = (int)cluster.num_blocks();
  const int rank = (int)cluster.block_rank();
  const int n = blockIdx.x / rs;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int ncols = nthr * PX;
  const int chunk = p.chunk;
  const int yb = rank * chunk, ye = min(p.iH - 1, yb + chunk - 1);

  float* rowA = sm;                                        // [MS][chunk] weight of row i towards output row dlo
  float* rowB = rowA + (size_t)MS * chunk;                 // [MS][chunk] ... towards dlo + 1
  int* rowD = reinterpret_cast<int*>(rowB + (size_t)MS * chunk);   // [MS][chunk] dlo
  int* fmask = rowD + (size_t)MS * chunk;                  // [chunk] bit s: scale s finishes an output row before row i
  float* wA = reinterpret_cast<float*>(fmask + chunk);     // [MS][ncols] x-weight towards output column dxl(x)
  float* wB = wA + (size_t)MS * ncols;                     // [MS][ncols] ... towards dxl(x) + 1
  int* colStart = reinterpret_cast<int*>(wB + (size_t)MS * ncols);   // per scale [oW + 1]
  float* totRow = reinterpret_cast<float*>(colStart + p.csTotal);    // per scale [maxRows] total vertical weight
  float* rowbuf = totRow + p.trTotal;                      // [NV][ncols] the parked output row
  float* part = rowbuf + (size_t)NV * ncols;               // per scale [maxRows][oW][NV + 1]

  TLD(0);
  // ---- tables ----
  for (int idx = tid; idx < MS * chunk; idx += nthr) {
    const int s = idx / chunk, i = idx - s * chunk, y = yb + i;
    float a = 0.f, b = 0.f;
    int d = 0;
    if (s < p.ns && y <= ye) {
      d = last_centre_le(y, p.hs[s], p.oH[s], p.iH);
      a = tri((float)y, src_pos(d, p.oH[s] - 1, p.iH - 1), p.hs[s]);
      b = d + 1 < p.oH[s] ? tri((float)y, src_pos(d + 1, p.oH[s] - 1, p.iH - 1), p.hs[s]) : 0.f;
    }
    rowA[idx] = a; rowB[idx] = b; rowD[idx] = d;
  }
  for (int idx = tid; idx < MAX_RANKS * MS; idx += nthr) {
    const int r = idx / MS, s = idx - r * MS;
    const int ybr = r * chunk, yer = min(p.iH - 1, ybr + chunk - 1);
    int f = 1, l = 0;
    if (r < rs && s < p.ns && ybr <= yer) {
      f = last_centre_le(ybr, p.hs[s], p.oH[s], p.iH);
      l = min(p.oH[s] - 1, last_centre_le(yer, p.hs[s], p.oH[s], p.iH) + 1);
      l = min(l, f + p.maxRows[s] - 1);
    }
    dF[r][s] = f; dL[r][s] = l;
  }
  for (int s = 0; s < p.ns; ++s) {
    const int oW = p.oW[s];
    int* cs = colStart + p.csOff[s];
    for (int dx = tid; dx <= oW; dx += nthr)
      cs[dx] = dx == 0 ? 0 : (dx == oW ? p.iW : (int)ceilf(src_pos(dx, oW - 1, p.iW - 1)));
    for (int x = tid; x < ncols; x += nthr) {
      const bool colok = x < p.iW;
      const int dxl = last_centre_le(min(x, p.iW - 1), p.ws[s], oW, p.iW);
      wA[s * ncols + x] = colok ? tri((float)x, src_pos(dxl, oW - 1, p.iW - 1), p.ws[s]) : 0.f;
      wB[s * ncols + x] = (colok && dxl + 1 < oW) ? tri((float)x, src_pos(dxl + 1, oW - 1, p.iW - 1), p.ws[s]) : 0.f;
    }
  }
  for (int i = tid; i < p.ptTotal; i += nthr) part[i] = 0.f;
  __syncthreads();
  for (int i = tid; i < chunk; i += nthr) {
    int m = 0;
    if (i > 0 && yb + i <= ye)
      for (int s = 0; s < p.ns; ++s) m |= (rowD[s * chunk + i] != rowD[s * chunk + i - 1]) ? (1 << s) : 0;
    fmask[i] = m;
  }
  for (int idx = tid; idx < p.trTotal; idx += nthr) {   // every (scale, local output row) on its own thread
      int s = 0;
      while (s + 1 < p.ns && p.trOff[s + 1] <= idx) ++s;
      const int dyl = idx - p.trOff[s];
      const int dy = dF[rank][s] + dyl;
      float t = 0.f;
      for (int i = 0; i <= ye - yb; ++i) {
        const int d = rowD[s * chunk + i];
        if (d == dy) t += rowA[s * chunk + i];
        if (d + 1 == dy) t += rowB[s * chunk + i];
      }
      totRow[p.trOff[s] + dyl] = t;
    }
  __syncthreads();

  TLD(1);
  // ---- horizontal pass of the parked row -> partial-result row dy of scale s ----
  auto horizontal = [&](int s, int dy) {
    __syncthreads();                                       // the parked row is complete
    const int dyl = dy - dF[rank][s];
    const int oW = p.oW[s];
    if (dyl >= 0 && dyl < p.maxRows[s] && dy < p.oH[s]) {
      const int* cs = colStart + p.csOff[s];
      const float* a_ = wA + s * ncols;
      const float* b_ = wB + s * ncols;
      int J = 1;
      while (J < 32 && J * 2 * oW <= nthr) J *= 2;
      const int j = tid % J;
      for (int it0 = 0; it0 < oW; it0 += nthr / J) {
        const int dx = it0 + tid / J;
        const bool live = dx < oW;
        float acc[NV];
#pragma unroll
        for (int k = 0; k < NV; ++k) acc[k] = 0.f;
        float wsum = 0.f;                                  // sum of the x-weights (for the total weight)
        if (live) {
          const int c1 = cs[dx], c2 = cs[dx + 1];
          if (dx > 0)
            for (int xx = cs[dx - 1] + j; xx < c1; xx += J) {   // columns whose right neighbour is dx
              const float w = b_[xx];
              wsum += w;
#pragma unroll
              for (int k = 0; k < NV; ++k) acc[k] = fmaf(w, rowbuf[k * ncols + xx], acc[k]);
            }
          for (int xx = c1 + j; xx < c2; xx += J) {              // columns whose left neighbour is dx
            const float w = a_[xx];
            wsum += w;
#pragma unroll
            for (int k = 0; k < NV; ++k) acc[k] = fmaf(w, rowbuf[k * ncols + xx], acc[k]);
          }
        }
        for (int off = J >> 1; off >= 1; off >>= 1) {
          wsum += __shfl_xor_sync(FULL, wsum, off);
#pragma unroll
          for (int k = 0; k < NV; ++k) acc[k] += __shfl_xor_sync(FULL, acc[k], off);
        }
        if (live && j == 0) {
          float* pr = part + p.ptOff[s] + ((size_t)dyl * oW + dx) * (NV + 1);
#pragma unroll
          for (int k = 0; k < NV; ++k) pr[k] = acc[k];
          pr[NV] = wsum * totRow[p.trOff[s] + dyl];
        }
      }
    }
    __syncthreads();                                       // the parked row may be overwritten
  };

  // ---- walk the rows ----
  float av[MS][2][NF], aw[MS][2][NF];
#pragma unroll
  for (int s = 0; s < MS; ++s)
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
      for (int f = 0; f < NF; ++f) av[s][k][f] = aw[s][k][f] = 0.f;
  const int x0 = tid * PX;
  const float* col = in + ((size_t)n * p.iH * p.iW + (size_t)min(x0, p.iW - PX)) * CT;

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
        const int m = fmask[i];
        if (m != 0) {                                      // uniform: some scale passed the centre of its next row
#pragma unroll
          for (int s = 0; s < MS; ++s)
            if ((m >> s) & 1) {
#pragma unroll
              for (int q = 0; q < PX; ++q)
#pragma unroll
                for (int c = 0; c < CT; ++c) {
                  rowbuf[c * ncols + x0 + q] = av[s][0][q * CT + c];
                  rowbuf[(CT + c) * ncols + x0 + q] = aw[s][0][q * CT + c];
                }
              horizontal(s, rowD[s * chunk + i - 1]);
#pragma unroll
              for (int f = 0; f < NF; ++f) {
                av[s][0][f] = av[s][1][f]; aw[s][0][f] = aw[s][1][f];
                av[s][1][f] = aw[s][1][f] = 0.f;
              }
            }
        }
        float sv[NF], ok[NF];
#pragma unroll
        for (int f = 0; f < NF; ++f) {                     // NaN samples only count in the total weight (:61-65)
          const bool fin = v[u][f] == v[u][f];
          sv[f] = fin ? v[u][f] : 0.f;
          ok[f] = fin ? 1.f : 0.f;
        }
#pragma unroll
        for (int s = 0; s < MS; ++s) {
          const float wa = rowA[s * chunk + i], wb = rowB[s * chunk + i];
#pragma unroll
          for (int f = 0; f < NF; ++f) {
            av[s][0][f] = fmaf(sv[f], wa, av[s][0][f]); aw[s][0][f] = fmaf(ok[f], wa, aw[s][0][f]);
            av[s][1][f] = fmaf(sv[f], wb, av[s][1][f]); aw[s][1][f] = fmaf(ok[f], wb, aw[s][1][f]);
          }
        }
      }
    }
  }
  TLD(2);
  if (yb <= ye) {
#pragma unroll
    for (int s = 0; s < MS; ++s)
      if (s < p.ns) {
        const int cur = rowD[s * chunk + (ye - yb)];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
#pragma unroll
          for (int q = 0; q < PX; ++q)
#pragma unroll
            for (int c = 0; c < CT; ++c) {
              rowbuf[c * ncols + x0 + q] = av[s][k][q * CT + c];
              rowbuf[(CT + c) * ncols + x0 + q] = aw[s][k][q * CT + c];
            }
          horizontal(s, cur + k);
        }
      }
  }

  TLD(3);
  // ---- combine the ranks' partial rows in rank order and finish (:72-76) ----
  cluster.sync();
  TLD(4);
  for (int s = 0; s < p.ns; ++s) {
    const int oW = p.oW[s];
    const int total = p.oH[s] * oW * CT;
    float* o = p.out[s] + (size_t)n * total;
    for (int idx = rank * nthr + tid; idx < total; idx += rs * nthr) {
      const int c = idx % CT, item = idx / CT;
      const int dy = item / oW, dx = item - dy * oW;
      float v_ = 0.f, w_ = 0.f, t_ = 0.f;
      for (int r = 0; r < rs; ++r) {
        if (dy < dF[r][s] || dy > dL[r][s]) continue;
        const float* rp = cluster.map_shared_rank(part, r) + p.ptOff[s] + ((size_t)(dy - dF[r][s]) * oW + dx) * (NV + 1);
        v_ += rp[c]; w_ += rp[CT + c]; t_ += rp[NV];
      }
      const float n_ = t_ - w_;                            // weight of the NaN samples (accum_nan)
      o[idx] = (n_ / w_ > 0.5f) ? nan_f() : v_ / w_;
    }
  }
  TLD(5);
  cluster.sync();   // keep this CTA's shared memory alive until every rank has read it
  TLD(6);
}

template <int CT, int PX, int U>
int launch_multi(const MsParams& p, int rs, int nthr, const float* in, cudaStream_t st) {
  auto kern = downsample_multi_kernel<CT, PX, U>;
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [&] {
    attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (attr_err == cudaSuccess) attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  });
  if (attr_err != cudaSuccess) { note_cuda_error(attr_err); return FN2_E_CUDA; }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)((long long)rs * p.N));
  cfg.blockDim = dim3((unsigned)nthr);
  cfg.dynamicSmemBytes = ms_smem_words(p.chunk, nthr * PX, 2 * CT, p.csTotal, p.trTotal, p.ptTotal) * sizeof(float);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)rs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p, in);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    if (rs > 8) return 2;   // a 16-CTA cluster does not fit this device / partition: the caller retries with 8
    note_cuda_error(e);
    return FN2_E_CUDA;
  }
  return after_launch();
}

// One fused launch for up to MS scales; returns 1 if this group of scales cannot take the fused kernel.
int try_fused(const float* in, float* const* outs, const int* out_hw, int ns, int N, int iH, int iW, int C,
              cudaStream_t st, int max_ranks) {
  if (C != 1 && C != 2) return 1;
  const int px = C == 1 ? (iW % 4 == 0 ? 4 : 1) : (iW % 2 == 0 ? 2 : 1);
  const int nthr = (((iW + px - 1) / px + 31) / 32) * 32;
  if (nthr > 512 || iW < px) return 1;
  if ((long long)N * MAX_RANKS > 0x7fffffffLL) return 1;
  int rs = 1;
  while (rs < max_ranks && iH >= 32 * rs) rs *= 2;          // >= 16 rows per rank
  MsParams p = {};
  p.N = N; p.iH = iH; p.iW = iW; p.ns = ns;
  p.chunk = (iH + rs - 1) / rs;
  int cs = 0, tr = 0, pt = 0;
  for (int s = 0; s < ns; ++s) {
    const int oH = out_hw[2 * s], oW = out_hw[2 * s + 1];
    p.oH[s] = oH; p.oW[s] = oW;
    p.ws[s] = (float)(iW - 1) / (float)(oW - 1);   // downsample_kernel_gpu.cu.cc:92-93
    p.hs[s] = (float)(iH - 1) / (float)(oH - 1);
    // centres at least 1.5 input pixels apart: an input row / column then feeds exactly two outputs per scale and
    // the walk never skips an output row, whatever the rounding of the centres
    if (!(p.hs[s] >= 1.5f) || !(p.ws[s] >= 1.5f) || oW > 1024) return 1;
    p.maxRows[s] = (int)((float)(p.chunk - 1) / p.hs[s]) + 3;
    p.csOff[s] = cs; p.trOff[s] = tr; p.ptOff[s] = pt;
    cs += oW + 1; tr += p.maxRows[s]; pt += p.maxRows[s] * oW * (2 * C + 1);
    p.out[s] = outs[s];
  }
  p.csTotal = cs; p.trTotal = tr; p.ptTotal = pt;
  if (ms_smem_words(p.chunk, nthr * px, 2 * C, cs, tr, pt) * sizeof(float) > 200 * 1024) return 1;
  if (C == 2) return px == 2 ? launch_multi<2, 2, 4>(p, rs, nthr, in, st) : launch_multi<2, 1, 4>(p, rs, nthr, in, st);
  return px == 4 ? launch_multi<1, 4, 4>(p, rs, nthr, in, st) : launch_multi<1, 1, 4>(p, rs, nthr, in, st);
}

}  // namespace

#ifdef FN2_TIMELINE
extern "C" __attribute__((visibility("default"))) int fn2_debug_timeline_ds(long long* host_out) {
  return (int)cudaMemcpyFromSymbol(host_out, g_tld, sizeof(long long) * 16);
}
#endif

int downsample_multi_forward(const float* in, float* const* outs, const int* out_hw, int num_scales, int N, int iH,
                             int iW, int C, cudaStream_t st) {
  for (int s0 = 0; s0 < num_scales; s0 += MS) {
    const int ns = num_scales - s0 < MS ? num_scales - s0 : MS;
    // 16-CTA clusters only up to ~1.3 CTAs per SM (measured on B200, 8 -> 16 ranks: N=4 0.079 -> 0.057 ms, N=12
    // 0.088 -> 0.067, but N=16 0.103 -> 0.134 and N=32 0.107 -> 0.212: 16 co-scheduled CTAs per GPC serialise once
    // the chip is full)
    static int big_ok = 1, sms = 0;
    if (sms == 0) {
      int dev = 0, v = 0;
      if (cudaGetDevice(&dev) == cudaSuccess) (void)cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
      sms = v > 0 ? v : 148;
    }
    int max_ranks = (big_ok && (long long)N * MAX_RANKS * 10 <= (long long)sms * 13) ? MAX_RANKS : 8;
    if (const char* e = getenv("FN2_DS_MAX_RANKS")) max_ranks = min(max(atoi(e), 1), MAX_RANKS);   // A/B switch
    int rc = try_fused(in, outs + s0, out_hw + 2 * s0, ns, N, iH, iW, C, st, max_ranks);
    if (rc == 2) {   // no room for 16-CTA clusters here: stay with the portable size from now on
      big_ok = 0;
      rc = try_fused(in, outs + s0, out_hw + 2 * s0, ns, N, iH, iW, C, st, 8);
    }
    if (rc == 1) {   // shapes the fused kernel does not take: one single-scale launch per size (downsample.cu)
      for (int s = s0; s < s0 + ns; ++s) {
        rc = downsample_forward(in, outs[s], N, iH, iW, C, out_hw[2 * s], out_hw[2 * s + 1], st);
        if (rc != FN2_OK) return rc;
      }
      rc = FN2_OK;
    }
    if (rc != FN2_OK) return rc;
  }
  return FN2_OK;
}

}  // namespace fn2
