// downsample_multi.cu -- all the loss scales of one tensor in ONE pass over the input, and (optionally) the loss terms
// that consume them.
//
// The reference's loss (flownet_c.py:127-166, flownet_s.py / flownet_sd.py likewise) scales the ground-truth flow
// (flow * 0.05), calls downsample(flow, size) five times on the same [N,384,512,2] tensor -- every call re-reading the
// whole input -- and reduces each scale against a prediction with average_endpoint_error (utils.py:209-224).  Each
// scale is the NaN-aware triangular-weighted average of downsample_kernel_gpu.cu.cc:35-76, and
// every scale is computed DIRECTLY from the input (the scale centres are not nested).
//
// Three launches, no atomics on data, a fixed summation order that does not depend on the batch size:
//   0. ms_tables_kernel   -- the weight tables, which depend only on the shapes: per (scale, input row) the output row
//      dlo below it and its two vertical weights, the same per input column, the column ranges of every output column,
//      the total weights, and which tiles feed which output.  One thread (or warp) per entry; the IEEE divisions of the
//      reference's formula are done once per launch here, not once per CTA.
//   1. ms_band_kernel     -- grid (row bands of 16 input rows) x (images): short CTAs, many waves.  A thread owns 4
//      adjacent floats of a row (16-byte loads, 4 rows in flight) and reads each exactly once.  The weight is separable,
//      w = wx(x) * wy(y), and an input row feeds at most two output rows of every scale (the last centre at or above
//      it, and the next one), so the thread keeps, per scale, the vertical value sums of those two live output rows in
//      registers (the NaN-weight sums live in thread-private shared-memory slots that are only touched once a warp
//      has met a NaN: none of that work is done for NaN-free input).  When the walk passes the centre of the next output row of a scale (a block-uniform
//      branch) the finished row is reduced horizontally BY EACH WARP ON ITS OWN: the lanes park their x-weighted column
//      sums in the warp's private shared-memory strip, and small lane groups add up the strip's share of every output
//      column it touches in a fixed order.  No block barrier, no cluster: warps never wait for each other.  The warp's
//      partial rows go to the caller's workspace (they stay in L2).
//   2. ms_combine_kernel  -- one thread per output pixel adds the partials of the (band, warp) tiles that feed it, in
//      tile order, applies the NaN rule (:72-76) and writes the downsampled value; when predictions are given it also
//      forms the end-point error sqrt(sum_c (pred - down)^2), the unit residual (pred - down) / epe that the loss
//      gradient needs, and the per-scale sums (block tree, then one ordered pass by the last block to finish).
// NaN samples: the reference keeps accum_weight (non-NaN samples) and accum_nan; here the NaN weight is accumulated
// itself (exactly zero when there is no NaN) and accum_weight = total weight - accum_nan, the total being a table.
#include <cstdlib>
#include "fn2_common.cuh"
#include "fn2_downsample.cuh"

namespace fn2 {
namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr int MS = 5;     // scales per launch (the reference's losses use five, flownet_c.py:135-160)
constexpr int RB = 16;    // input rows per band: fixed, so the summation order never depends on the batch size
constexpr int U = 4;      // rows in flight per thread
constexpr int CB = 256;   // threads of the combine kernel
constexpr int NSLOT = 2 * MS;   // staged rows: two groups of MS used alternately, all of them by the band's last flush
#ifndef DS_MINB
#define DS_MINB 2
#endif
static_assert(RB * 4 <= 64, "one NaN bit per (row, float) of a thread in a 64-bit word");

// Offsets (4-byte words) of the per-scale tables inside the workspace.
struct MsTables {
  int rowW[MS];    // [iH] float2: weight of input row y towards output rows dlo(y), dlo(y) + 1
  int rowD[MS];    // [iH] int: dlo(y)
  int colW[MS];    // [iW] float2: the same along x
  int colS[MS];    // [oW + 1] int: first input column whose nearest-left output column is dx
  int totX[MS];    // [oW] float: sum of the x-weights of output column dx
  int totY[MS];    // [oH] float
  int bandR[MS];   // [oH] int2: first / last band that feeds output row dy
  int words;       // size of the table area
};

struct MsParams {
  int N, iH, iW, ns;
  int nbands, nthr;               // bands per image, threads of a band CTA (a multiple of 32)
  int oH[MS], oW[MS];
  int maxRows[MS];                // output rows a band can touch
  int J[MS];                      // lanes that share one output column in the horizontal pass (a power of two <= 32)
  int pOff[MS], pTile;            // partial entries: offset of scale s inside a band's tile, entries per tile
  int blkOff[MS + 1];             // combine kernel: first block of scale s within an image
  float hs[MS], ws[MS];
  float in_scale;                 // samples are multiplied by this on load (the loss's flow * 0.05), 1 = none
  MsTables t;
  float* out[MS];                 // downsampled tensors (may be null when only the loss terms are wanted)
  const float* pred[MS];          // predictions to compare with (null: plain downsample)
  float* dir[MS];                 // (pred - down) / epe per element (may be null)
  float* epe;                     // [ns] per-scale sum of end-point errors, divided by N (null without pred)
  size_t ctrlOff, partOff;        // byte offsets in the workspace: control block (ticket + block sums), partials
};

// Shared-memory layout of the band kernel (bytes, every area 16-byte aligned); used by the host for the launch size.
struct BandSmem {
  size_t rowW, rowD, fmask, colS, colW, stage, nstage, total;
};
__host__ __device__ inline size_t up16(size_t v) { return (v + 15) & ~(size_t)15; }
__host__ __device__ inline BandSmem band_smem_layout(int nthr, int iW, int cswords) {
  BandSmem L;
  size_t o = 0;
  L.rowW = o; o = up16(o + (size_t)MS * RB * 8);
  L.rowD = o; o = up16(o + (size_t)MS * RB * 4);
  L.fmask = o; o = up16(o + (size_t)RB * 4);
  L.colS = o; o = up16(o + (size_t)cswords * 4);
  L.colW = o; o = up16(o + (size_t)MS * iW * 8);
  L.stage = o; o = up16(o + (size_t)NSLOT * nthr * 16);
  L.nstage = o; o = up16(o + (size_t)nthr * 16);
  L.total = o;
  return L;
}

// ---------------------------------------------------------------------------------------------------- tables
// grid (TBX, ns): blockIdx.y = scale; the entries of a scale are independent, each on its own thread (rows, columns,
// ranges) or its own warp (total weights), so the kernel is as long as ONE entry's chain of IEEE divisions.
constexpr int TBX = 24;
__global__ void __launch_bounds__(256) ms_tables_kernel(MsParams p, float* __restrict__ ws) {
  int* wi = reinterpret_cast<int*>(ws);
  const int s = blockIdx.y;
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int gthreads = gridDim.x * blockDim.x;
  const int gwarp = gtid >> 5, gwarps = gthreads >> 5, lane = threadIdx.x & 31;
  if (gtid == 0 && s == 0) *reinterpret_cast<unsigned*>(reinterpret_cast<char*>(ws) + p.ctrlOff) = 0u;   // combine ticket
  const int oH = p.oH[s], oW = p.oW[s];
  const float hs = p.hs[s], wsc = p.ws[s];
  // task space: [0, iH) input rows, then iW input columns, then oW + 1 column starts, then oH output rows
  const int ntask = p.iH + p.iW + (oW + 1) + oH;
  for (int task = gtid; task < ntask; task += gthreads) {
    if (task < p.iH) {
      const int y = task;
      const int d = last_centre_le(y, hs, oH, p.iH);
      const float a = tri((float)y, src_pos(d, oH - 1, p.iH - 1), hs);
      const float b = d + 1 < oH ? tri((float)y, src_pos(d + 1, oH - 1, p.iH - 1), hs) : 0.f;
      reinterpret_cast<float2*>(ws + p.t.rowW[s])[y] = make_float2(a, b);
      wi[p.t.rowD[s] + y] = d;
    } else if (task < p.iH + p.iW) {
      const int x = task - p.iH;
      const int d = last_centre_le(x, wsc, oW, p.iW);
      const float a = tri((float)x, src_pos(d, oW - 1, p.iW - 1), wsc);
      const float b = d + 1 < oW ? tri((float)x, src_pos(d + 1, oW - 1, p.iW - 1), wsc) : 0.f;
      reinterpret_cast<float2*>(ws + p.t.colW[s])[x] = make_float2(a, b);
    } else if (task < p.iH + p.iW + oW + 1) {
      const int dx = task - p.iH - p.iW;
      wi[p.t.colS[s] + dx] = dx == 0 ? 0 : (dx == oW ? p.iW : (int)ceilf(src_pos(dx, oW - 1, p.iW - 1)));
    } else {
      // band b feeds dy iff it holds a row y with dlo(y) in {dy - 1, dy}   (dlo(y) >= d  <=>  y >= centre(d))
      const int dy = task - p.iH - p.iW - oW - 1;
      const int ylo = dy >= 2 ? (int)ceilf(src_pos(dy - 1, oH - 1, p.iH - 1)) : 0;
      const int yhi = dy + 1 < oH ? (int)ceilf(src_pos(dy + 1, oH - 1, p.iH - 1)) - 1 : p.iH - 1;
      reinterpret_cast<int2*>(wi + p.t.bandR[s])[dy] =
          make_int2(min(ylo, p.iH - 1) / RB, min(max(yhi, 0), p.iH - 1) / RB);
    }
  }
  // one warp per output row / column: total weights (lane-strided, then a fixed tree); warps taken from the END of the
  // grid so they are not the ones busy with the entries above
  for (int e = gwarps - 1 - gwarp; e < oH + oW; e += gwarps) {
    const bool isx = e >= oH;
    const int d = isx ? e - oH : e;
    const int on = isx ? oW : oH, in_ = isx ? p.iW : p.iH;
    const float sc = isx ? wsc : hs;
    const float c = src_pos(d, on - 1, in_ - 1);
    const int lo = max(0, (int)floorf(c - sc)), hi = min(in_ - 1, (int)ceilf(c + sc));
    float t = 0.f;
    for (int i = lo + lane; i <= hi; i += 32) t += tri((float)i, c, sc);
    for (int off = 16; off >= 1; off >>= 1) t += __shfl_xor_sync(FULL, t, off);
    if (lane == 0) ws[(isx ? p.t.totX[s] : p.t.totY[s]) + d] = t;
  }
}

template <int NV>
__device__ __forceinline__ void add_entry(float (&acc)[NV], const float* e) {
  if constexpr (NV == 4) {
    const float4 t = *reinterpret_cast<const float4*>(e);
    acc[0] += t.x; acc[1] += t.y; acc[2] += t.z; acc[3] += t.w;
  } else {
    const float2 t = *reinterpret_cast<const float2*>(e);
    acc[0] += t.x; acc[1] += t.y;
  }
}

// ------------------------------------------------------------------------------------------------- band walk
// grid (bands of RB input rows) x (images); CT channels; a thread owns the 4 floats (PX = 4 / CT adjacent pixels) of one
// 16-byte column slot and walks the band's rows downwards, U rows in flight, reading every input byte exactly once.
// The weight is separable, w = wx(x) * wy(y), and an input row feeds at most two output rows of every scale (the last
// centre at or above it, and the next one): per scale the thread keeps the VERTICAL sums of those two live output rows
// of its 4 floats in registers (2 FMAs per float, scale and row).  When the walk passes the centre of the next output
// row of some scales (a block-uniform event, fmask) the finished rows are parked in shared memory -- one 16-byte store
// per thread and row --, ONE block barrier follows, and the whole CTA sums each parked row horizontally: J lanes share
// an output column, each adds its share of the column's window in a fixed order (x-weights from a shared copy of the
// table), a fixed shuffle tree joins the J partial sums, and the band's partial output row goes to the caller's
// workspace (it stays in L2).  Parked rows alternate between two groups of slots, so no second barrier is needed.
// NaN samples are rare: the walk only records WHERE they were (one bit per row and float); the barrier of a flush also
// tells whether any thread of the CTA has met one (__syncthreads_or), and only then the NaN weights of the parked rows
// are rebuilt from the bits (same additions, same order as a running sum) and summed like the values.
template <int CT, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB) ms_band_kernel(MsParams p, const float* __restrict__ in,
                                                             float* __restrict__ ws) {
  constexpr int PX = 4 / CT;
  constexpr int NV = 2 * CT;                    // per output pixel: value and NaN-weight sums per channel
  extern __shared__ __align__(16) uint8_t sm[];
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int lane = tid & 31;
  const int band = blockIdx.x, n = blockIdx.y;
  const int yb = band * RB, ye = min(p.iH - 1, yb + RB - 1);
  const int nrows = ye - yb + 1;
  const int iW = p.iW;
  const int* wsi = reinterpret_cast<const int*>(ws);

  // the colS tables are packed back to back in the workspace and followed by totX[0]
  const int cswords = p.t.totX[0] - p.t.colS[0];
  const BandSmem L = band_smem_layout(nthr, iW, cswords);
  float2* rowW = reinterpret_cast<float2*>(sm + L.rowW);     // [MS][RB]
  int* rowD = reinterpret_cast<int*>(sm + L.rowD);           // [MS][RB]
  int* fmask = reinterpret_cast<int*>(sm + L.fmask);         // [RB] bit s: scale s finishes an output row before row i
  int* colS = reinterpret_cast<int*>(sm + L.colS);           // per scale [oW + 1], packed
  float2* colW = reinterpret_cast<float2*>(sm + L.colW);     // [MS][iW]
  float4* stage = reinterpret_cast<float4*>(sm + L.stage);   // [NSLOT][nthr]: parked rows of vertical sums
  float4* nstage = reinterpret_cast<float4*>(sm + L.nstage); // [nthr]: NaN weights of one parked row (rare path)

  // the first rows are requested before anything else so that their latency overlaps the table copy
  const int x0 = tid * PX;
  const float* col = in + ((size_t)n * p.iH * iW + (size_t)min(x0, iW - PX)) * CT;
  float4 v[U];
#pragma unroll
  for (int u = 0; u < U; ++u) v[u] = __ldg(reinterpret_cast<const float4*>(col + (size_t)min(yb + u, ye) * iW * CT));

  // ---- this band's slice of the row tables, the column tables ----
  for (int idx = tid; idx < MS * RB; idx += nthr) {
    const int s = idx / RB, i = idx - s * RB;
    float2 w = make_float2(0.f, 0.f);
    int d = 0;
    if (s < p.ns && i < nrows) {
      w = __ldg(reinterpret_cast<const float2*>(ws + p.t.rowW[s]) + yb + i);
      d = __ldg(wsi + p.t.rowD[s] + yb + i);
    }
    rowW[idx] = w; rowD[idx] = d;
  }
  for (int s = 0; s < p.ns; ++s)
    for (int x = tid; x < iW; x += nthr) colW[s * iW + x] = __ldg(reinterpret_cast<const float2*>(ws + p.t.colW[s]) + x);
  for (int i = tid; i < cswords; i += nthr) colS[i] = __ldg(wsi + p.t.colS[0] + i);
  __syncthreads();
  for (int i = tid; i < RB; i += nthr) {
    int m = 0;
    if (i > 0 && i < nrows)
      for (int s = 0; s < p.ns; ++s) m |= (rowD[s * RB + i] != rowD[s * RB + i - 1]) ? (1 << s) : 0;
    fmask[i] = m;
  }
  __syncthreads();

  float* part = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + p.partOff) +
                ((size_t)n * p.nbands + band) * p.pTile * NV;

  // ---- horizontal pass of one parked row (scale s, output row dy), by the whole CTA ----
  auto hpass = [&](int s, int dy, int slot, bool withnan) {
    const float* st = reinterpret_cast<const float*>(stage + (size_t)slot * nthr);
    const float* nst = reinterpret_cast<const float*>(nstage);
    const int oW = p.oW[s];
    const int J = p.J[s], cpp = nthr / J;
    const int* cs = colS + (p.t.colS[s] - p.t.colS[0]);
    const float2* cw = colW + s * iW;
    float* prow = part + ((size_t)p.pOff[s] + (size_t)(dy - rowD[s * RB]) * oW) * NV;
    const int j = tid & (J - 1), c = tid / J;
    for (int c0 = 0; c0 < oW; c0 += cpp) {
      const int dx = c0 + c;
      const bool live = dx < oW;
      float acc[NV];
#pragma unroll
      for (int k = 0; k < NV; ++k) acc[k] = 0.f;
      if (live) {
        const int xa = dx > 0 ? cs[dx - 1] : 0, xm = cs[dx], xb = cs[dx + 1];
        for (int x = xa + j; x < xb; x += J) {      // columns whose right (x < xm) or left neighbour is dx
          const float2 w = cw[x];
          const float wt = x < xm ? w.y : w.x;
          if constexpr (CT == 2) {
            const float2 a = reinterpret_cast<const float2*>(st)[x];
            acc[0] = fmaf(wt, a.x, acc[0]); acc[1] = fmaf(wt, a.y, acc[1]);
            if (withnan) {
              const float2 b = reinterpret_cast<const float2*>(nst)[x];
              acc[2] = fmaf(wt, b.x, acc[2]); acc[3] = fmaf(wt, b.y, acc[3]);
            }
          } else {
            acc[0] = fmaf(wt, st[x], acc[0]);
            if (withnan) acc[1] = fmaf(wt, nst[x], acc[1]);
          }
        }
      }
      for (int off = J >> 1; off >= 1; off >>= 1) {
#pragma unroll
        for (int k = 0; k < NV; ++k) acc[k] += __shfl_xor_sync(FULL, acc[k], off);
      }
      if (live && j == 0) {
        if constexpr (NV == 4) *reinterpret_cast<float4*>(prow + dx * 4) = make_float4(acc[0], acc[1], acc[2], acc[3]);
        else *reinterpret_cast<float2*>(prow + dx * 2) = make_float2(acc[0], acc[1]);
      }
    }
  };

  uint64_t nanbits = 0;       // bit 4 * i + f: float f of this thread was NaN in row i of the band
  // NaN weight of this thread's floats towards output row dy of scale s: the running sum the walk would have kept
  auto nan_weights = [&](int s, int dy) {
    float a[4] = {0.f, 0.f, 0.f, 0.f};
    if (nanbits != 0) {
      for (int i = 0; i < nrows; ++i) {
        const unsigned bits = (unsigned)(nanbits >> (4 * i)) & 15u;
        if (bits == 0) continue;
        const int d = rowD[s * RB + i];
        const float2 w = rowW[s * RB + i];
        const float wt = d == dy ? w.x : (d + 1 == dy ? w.y : 0.f);
#pragma unroll
        for (int f = 0; f < 4; ++f) a[f] = fmaf((bits >> f) & 1u ? 1.f : 0.f, wt, a[f]);
      }
    }
    return make_float4(a[0], a[1], a[2], a[3]);
  };

  // flush of the items (scale s, live row k) whose bit 2 * s + k is set in `items`; `group` picks the slots
  auto flush = [&](unsigned items, int group, const float (&av)[MS][2][4], int i_last) {
    int slot = group * MS;
#pragma unroll
    for (int s = 0; s < MS; ++s)
#pragma unroll
      for (int k = 0; k < 2; ++k)
        if ((items >> (2 * s + k)) & 1u)
          stage[(size_t)(slot++) * nthr + tid] = make_float4(av[s][k][0], av[s][k][1], av[s][k][2], av[s][k][3]);
    const bool any = __syncthreads_or(nanbits != 0);
    slot = group * MS;
    for (int s = 0; s < MS; ++s)
      for (int k = 0; k < 2; ++k)
        if ((items >> (2 * s + k)) & 1u) {
          const int dy = rowD[s * RB + i_last] + k;
          if (dy < p.oH[s]) {                                // uniform
            if (!any) {
              hpass(s, dy, slot, false);
            } else {
              nstage[tid] = nan_weights(s, dy);
              __syncthreads();
              hpass(s, dy, slot, true);
              __syncthreads();
            }
          }
          ++slot;
        }
  };

  // ---- walk the rows ----
  float av[MS][2][4];
#pragma unroll
  for (int s = 0; s < MS; ++s)
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
      for (int f = 0; f < 4; ++f) av[s][k][f] = 0.f;
  const float scale = p.in_scale;
  int group = 0;

  for (int y0 = yb; y0 <= ye; y0 += U) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const float4 cur = v[u];
      if (y0 + U + u <= ye)                                // uniform: U loads stay in flight
        v[u] = __ldg(reinterpret_cast<const float4*>(col + (size_t)(y0 + U + u) * iW * CT));
      if (y0 + u <= ye) {                                  // uniform
        const int i = y0 + u - yb;
        const int m = fmask[i];
        if (m != 0) {                                      // uniform: some scales passed the centre of their next row
          unsigned items = 0;
#pragma unroll
          for (int s = 0; s < MS; ++s) items |= ((m >> s) & 1) ? (1u << (2 * s)) : 0u;
          flush(items, group, av, i - 1);
          group ^= 1;
#pragma unroll
          for (int s = 0; s < MS; ++s)
            if ((m >> s) & 1) {
#pragma unroll
              for (int f = 0; f < 4; ++f) { av[s][0][f] = av[s][1][f]; av[s][1][f] = 0.f; }
            }
        }
        // flow * in_scale is a rounded fp32 product, as if the scaled tensor had been materialised (flownet_c.py:128)
        const float x[4] = {__fmul_rn(cur.x, scale), __fmul_rn(cur.y, scale), __fmul_rn(cur.z, scale),
                            __fmul_rn(cur.w, scale)};
        float sv[4];
#pragma unroll
        for (int f = 0; f < 4; ++f) {                      // NaN samples only count in the NaN weight (:61-65)
          const bool fin = x[f] == x[f];
          sv[f] = fin ? x[f] : 0.f;
          if (!fin) nanbits |= 1ull << (4 * i + f);
        }
#pragma unroll
        for (int s = 0; s < MS; ++s) {
          const float2 w = rowW[s * RB + i];
#pragma unroll
          for (int f = 0; f < 4; ++f) {
            av[s][0][f] = fmaf(sv[f], w.x, av[s][0][f]);
            av[s][1][f] = fmaf(sv[f], w.y, av[s][1][f]);
          }
        }
      }
    }
  }
  // the band's last flush parks both live rows of every scale: it uses both slot groups, so every thread must have
  // left the previous flush's horizontal pass
  __syncthreads();
  flush((1u << (2 * p.ns)) - 1u, 0, av, nrows - 1);
}

// --------------------------------------------------------------------------------------------------- combine
template <int CT>
__global__ void __launch_bounds__(CB) ms_combine_kernel(MsParams p, float* __restrict__ ws) {
  constexpr int NV = 2 * CT;
  __shared__ float red[CB / 32];
  __shared__ bool lastBlock;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int bpi = p.blkOff[p.ns];
  const int n = blockIdx.x / bpi, rb = blockIdx.x - n * bpi;
  int s = 0;
  while (s + 1 < p.ns && p.blkOff[s + 1] <= rb) ++s;
  const int oH = p.oH[s], oW = p.oW[s];
  const int idx = (rb - p.blkOff[s]) * CB + tid;
  const bool live = idx < oH * oW;
  const int* wsi = reinterpret_cast<const int*>(ws);
  const float* parts = reinterpret_cast<const float*>(reinterpret_cast<const char*>(ws) + p.partOff);
  float epe = 0.f;
  if (live) {
    const int dy = idx / oW, dx = idx - dy * oW;
    const int2 br = __ldg(reinterpret_cast<const int2*>(wsi + p.t.bandR[s]) + dy);
    float acc[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) acc[k] = 0.f;
    for (int b = br.x; b <= br.y; ++b) {                     // the bands that feed this row, in band order
      const int slot = dy - __ldg(wsi + p.t.rowD[s] + b * RB);
      const float* e = parts + ((((size_t)n * p.nbands + b) * p.pTile + p.pOff[s] + (size_t)slot * oW + dx)) * NV;
      add_entry<NV>(acc, e);
    }
    const float tot = __ldg(ws + p.t.totY[s] + dy) * __ldg(ws + p.t.totX[s] + dx);
    const size_t o = (((size_t)n * oH + dy) * oW + dx) * CT;
    float e2 = 0.f, diff[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) {
      const float nanw = acc[CT + c];                        // accum_nan
      const float wgt = tot - nanw;                          // accum_weight: the samples that are not NaN
      // :72-76.  accum_nan / accum_weight > 0.5 means accum_weight < 2/3 of the total, far from the cancellation in
      // `tot - nanw`; when (nearly) every sample is NaN the difference is rounding noise around 0 (the reference has
      // an exact 0 there and divides by it): <= 0 is NaN too.
      const float r = (wgt <= 0.f || nanw / wgt > 0.5f) ? nan_f() : acc[c] / wgt;
      if (p.out[s]) p.out[s][o + c] = r;
      if (p.pred[s]) {
        diff[c] = p.pred[s][o + c] - r;
        e2 = fmaf(diff[c], diff[c], e2);
      }
    }
    if (p.pred[s]) {
      epe = sqrtf(e2);                                       // utils.py:219-223
      if (p.dir[s]) {
#pragma unroll
        for (int c = 0; c < CT; ++c) p.dir[s][o + c] = diff[c] / epe;   // 0/0 = NaN, as the gradient of sqrt at 0
      }
    }
  }
  if (p.epe == nullptr) return;                              // uniform: plain downsample
  // ---- per-scale sums: block tree, then the last block to finish adds the block sums in block order ----
  for (int off = 16; off >= 1; off >>= 1) epe += __shfl_xor_sync(FULL, epe, off);
  if (lane == 0) red[warp] = epe;
  __syncthreads();
  char* ctrl = reinterpret_cast<char*>(ws) + p.ctrlOff;
  unsigned* ticket = reinterpret_cast<unsigned*>(ctrl);
  float* bsum = reinterpret_cast<float*>(ctrl + 16);
  if (tid == 0) {
    float t = 0.f;
    for (int i = 0; i < CB / 32; ++i) t += red[i];
    bsum[blockIdx.x] = t;
    __threadfence();
    lastBlock = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!lastBlock) return;
  __threadfence();
  if (warp < p.ns) {                                         // one warp per scale, lane-strided then a fixed tree
    const int sc = warp;
    const int nb = p.blkOff[sc + 1] - p.blkOff[sc];
    float t = 0.f;
    for (int i = lane; i < p.N * nb; i += 32) t += __ldcg(bsum + (size_t)(i / nb) * bpi + p.blkOff[sc] + i % nb);
    for (int off = 16; off >= 1; off >>= 1) t += __shfl_xor_sync(FULL, t, off);
    if (lane == 0) p.epe[sc] = t / (float)p.N;              // utils.py:224: reduce_sum(epe) / num_samples
  }
  if (tid == 0) *ticket = 0u;
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Fills the shape-dependent part of the parameters; returns false if these scales cannot take the fused kernels.
bool ms_plan(MsParams* pp, const int* out_hw, int ns, int N, int iH, int iW, int C, size_t* ws_bytes,
             size_t* band_smem) {
  MsParams& p = *pp;
  if (C != 1 && C != 2) return false;
  const int px = 4 / C;
  if (iW % px != 0 || iW < px) return false;
  const int nthr = (((iW / px) + 31) / 32) * 32;
  if (nthr > 512) return false;
  if (ns < 1 || ns > MS || N > 65535) return false;
  p.N = N; p.iH = iH; p.iW = iW; p.ns = ns;
  p.nbands = (iH + RB - 1) / RB;
  p.nthr = nthr;
  int words = 0, pt = 0, blk = 0;
  auto take = [&](int n) { const int at = words; words += (n + 3) / 4 * 4; return at; };   // 16-byte aligned tables
  for (int s = 0; s < ns; ++s) {
    const int oH = out_hw[2 * s], oW = out_hw[2 * s + 1];
    p.oH[s] = oH; p.oW[s] = oW;
    p.ws[s] = (float)(iW - 1) / (float)(oW - 1);   // downsample_kernel_gpu.cu.cc:92-93
    p.hs[s] = (float)(iH - 1) / (float)(oH - 1);
    // centres at least 1.5 input pixels apart: an input row / column then feeds exactly two outputs per scale and
    // the walk never skips an output row, whatever the rounding of the centres
    if (!(p.hs[s] >= 1.5f) || !(p.ws[s] >= 1.5f)) return false;
    p.maxRows[s] = (int)((float)(RB - 1) / p.hs[s]) + 3;
    int J = 1;   // the CTA's threads are dealt out J to an output column
    while (J < 32 && J * 2 * oW <= nthr) J *= 2;
    p.J[s] = J;
    p.pOff[s] = pt;
    pt += p.maxRows[s] * oW;
    p.blkOff[s] = blk;
    blk += (oH * oW + CB - 1) / CB;
  }
  for (int s = ns; s < MS; ++s) {   // unused scales: harmless values (their weights are zero, they are never flushed)
    p.oH[s] = p.oW[s] = 2; p.J[s] = 1; p.maxRows[s] = 0; p.pOff[s] = pt; p.hs[s] = p.ws[s] = 2.f;
  }
  for (int s = ns; s <= MS; ++s) p.blkOff[s] = blk;
  p.pTile = pt;
  // table area; the colS tables are packed back to back and followed by totX[0] (the band kernel copies them as one run)
  for (int s = 0; s < ns; ++s) { p.t.rowW[s] = take(2 * iH); p.t.rowD[s] = take(iH); }
  for (int s = 0; s < ns; ++s) p.t.colW[s] = take(2 * iW);
  for (int s = 0; s < ns; ++s) { p.t.colS[s] = words; words += p.oW[s] + 1; }
  words = (words + 3) / 4 * 4;
  for (int s = 0; s < ns; ++s) p.t.totX[s] = take(p.oW[s]);
  for (int s = 0; s < ns; ++s) p.t.totY[s] = take(p.oH[s]);
  for (int s = 0; s < ns; ++s) p.t.bandR[s] = take(2 * p.oH[s]);
  for (int s = ns; s < MS; ++s)
    p.t.rowW[s] = p.t.rowD[s] = p.t.colW[s] = p.t.colS[s] = p.t.totX[s] = p.t.totY[s] = p.t.bandR[s] = 0;
  p.t.words = words;
  const long long nblocks = (long long)N * blk;
  if (nblocks > 0x7fffffffLL) return false;
  p.ctrlOff = align_up((size_t)words * 4, 256);
  p.partOff = align_up(p.ctrlOff + 16 + (size_t)nblocks * 4, 256);
  *ws_bytes = p.partOff + (size_t)N * p.nbands * p.pTile * (2 * C) * sizeof(float);
  *band_smem = band_smem_layout(nthr, iW, p.t.totX[0] - p.t.colS[0]).total;
  return *band_smem <= 200 * 1024;
}

template <int CT>
int ms_launch(const MsParams& p, size_t band_smem, const float* in, void* ws, cudaStream_t st) {
  static PerDeviceOnce once;   // per template instantiation; function attributes are per device
  const cudaError_t attr_err = per_device_once(once, [] {
    cudaError_t e = cudaFuncSetAttribute(ms_band_kernel<CT, 256, DS_MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         200 * 1024);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(ms_band_kernel<CT, 512, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    return e;
  });
  if (attr_err != cudaSuccess) { note_cuda_error(attr_err); (void)cudaGetLastError(); return FN2_E_CUDA; }
  float* w = static_cast<float*>(ws);
  ms_tables_kernel<<<dim3(TBX, (unsigned)p.ns), 256, 0, st>>>(p, w);
  int rc = after_launch();
  if (rc != FN2_OK) return rc;
  const dim3 grid((unsigned)p.nbands, (unsigned)p.N);
  if (p.nthr <= 256)
    ms_band_kernel<CT, 256, DS_MINB><<<grid, p.nthr, band_smem, st>>>(p, in, w);
  else
    ms_band_kernel<CT, 512, 1><<<grid, p.nthr, band_smem, st>>>(p, in, w);
  rc = after_launch();
  if (rc != FN2_OK) return rc;
  ms_combine_kernel<CT><<<(unsigned)(p.N * p.blkOff[p.ns]), CB, 0, st>>>(p, w);
  return after_launch();
}

}  // namespace

size_t multiscale_workspace_bytes(const int* out_hw, int num_scales, int N, int iH, int iW, int C) {
  size_t best = 0;
  for (int s0 = 0; s0 < num_scales; s0 += MS) {
    const int ns = num_scales - s0 < MS ? num_scales - s0 : MS;
    MsParams p = {};
    size_t wsb = 0, smem = 0;
    if (ms_plan(&p, out_hw + 2 * s0, ns, N, iH, iW, C, &wsb, &smem) && wsb > best) best = wsb;
  }
  return best;
}

// predictions / residual_dir / epe_sums may be null (plain multi-scale downsample); outs entries may be null when
// predictions are given (only the loss terms wanted).
int multiscale_forward(const float* in, float in_scale, float* const* outs, const float* const* preds,
                       float* const* dirs, float* epe_sums, const int* out_hw, int num_scales, int N, int iH, int iW,
                       int C, void* ws, size_t ws_bytes, cudaStream_t st) {
  for (int s0 = 0; s0 < num_scales; s0 += MS) {
    const int ns = num_scales - s0 < MS ? num_scales - s0 : MS;
    MsParams p = {};
    size_t need = 0, smem = 0;
    if (!ms_plan(&p, out_hw + 2 * s0, ns, N, iH, iW, C, &need, &smem)) {
      // shapes the fused kernels do not take: one single-scale launch per size (downsample.cu); the loss terms need
      // the fused path
      if (preds) return FN2_E_UNSUPPORTED_GEOMETRY;
      if (in_scale != 1.0f) return FN2_E_UNSUPPORTED_GEOMETRY;
      for (int s = s0; s < s0 + ns; ++s) {
        const int rc = downsample_forward(in, outs[s], N, iH, iW, C, out_hw[2 * s], out_hw[2 * s + 1], st);
        if (rc != FN2_OK) return rc;
      }
      continue;
    }
    if (ws_bytes < need || !ws) return FN2_E_WORKSPACE;
    p.in_scale = in_scale;
    for (int s = 0; s < ns; ++s) {
      p.out[s] = outs ? outs[s0 + s] : nullptr;
      p.pred[s] = preds ? preds[s0 + s] : nullptr;
      p.dir[s] = dirs ? dirs[s0 + s] : nullptr;
    }
    p.epe = preds ? epe_sums + s0 : nullptr;
    const int rc = C == 2 ? ms_launch<2>(p, smem, in, ws, st) : ms_launch<1>(p, smem, in, ws, st);
    if (rc != FN2_OK) return rc;
  }
  return FN2_OK;
}

}  // namespace fn2
