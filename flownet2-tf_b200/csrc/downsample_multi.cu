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
//      dlo below it and its two vertical weights, the same per input column, the first input column of every output
//      column's interval, the total weights, and which bands feed which output row.  One thread (or warp) per entry;
//      the IEEE divisions of the reference's formula are done once per launch here, not once per CTA.
//   1. ms_band_kernel     -- grid (bands of 16 input rows) x (images).  The band (64 KB for a 512 x 2 flow) is copied
//      into shared memory by the bulk-copy engine in one go; nothing else ever reads the input.  The weight is
//      separable, w = wx(x) * wy(y).  HORIZONTAL FIRST: the input columns between two output-column centres of a scale
//      form an interval whose pixels feed exactly those two output columns; J lanes share an interval, each owning a
//      16-byte-aligned run of at most 10 pixels whose x-weights (towards the left centre, A, and the right one, B) sit
//      in registers for the whole band.  Per input row a thread reads its run once (<= 5 16-byte loads, conflict
//      free), forms the two x-weighted sums, and adds them, y-weighted, to the vertical sums of the (at most) two
//      live output rows of its scale -- all in registers, no barrier, no shuffle.  When the walk passes the centre
//      of the next output row of the thread's scale (warp-uniform: a warp serves one scale) the J lanes join their
//      sums with a fixed shuffle tree and one lane stores the partial (A and B part of the interval) to the caller's
//      workspace (it stays in L2).  Every pixel is read from shared memory once per scale.
//   2. ms_combine_kernel  -- one thread per output pixel adds, band by band, the A part of its own interval and the B
//      part of the interval to its left, applies the NaN rule (:72-76) and writes the downsampled value; when
//      predictions are given it also forms the end-point error sqrt(sum_c (pred - down)^2), the unit residual
//      (pred - down) / epe that the loss gradient needs, and the per-scale sums (block tree, then one ordered pass by
//      the last block to finish).
// NaN samples: the fast path multiplies and adds blindly; a NaN (or an infinity meeting a zero weight) turns the row's
// sums into NaN, which sends THAT thread and row through the careful path (sample != sample -> counted in the NaN
// weight instead, :61-65) and marks the row; the NaN weights are rebuilt from the marked rows (still in shared memory)
// when an output row is stored.  accum_weight = total weight - accum_nan, the total being a table.
#include <cstdlib>
#include "fn2_common.cuh"
#include "fn2_downsample.cuh"

namespace fn2 {
namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr int MS = 5;     // scales per launch (the reference's losses use five, flownet_c.py:135-160)
constexpr int RB = 16;    // input rows per band: fixed, so the summation order never depends on the batch size
constexpr int CB = 256;   // threads of the combine kernel
constexpr int MAXBANDS = 1024;   // bands per image the combine kernel keeps a table for (iH <= 16384)
static_assert(RB <= 32, "one marked-row bit per row of the band");

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
  int J[MS], L[MS];               // lanes per interval (a power of two <= 32), pixels per lane (a multiple of 4 / C)
  int tOff[MS + 1];               // first thread of scale s in a band CTA (multiples of 32)
  int pOff[MS], pTile;            // partial entries: offset of scale s inside a band's tile, entries per tile
  int blkOff[MS + 1];             // combine kernel: first block of scale s within an image
  float hs[MS], ws[MS];
  float in_scale;                 // samples are multiplied by this on load (the loss's flow * 0.05), 1 = none
  MsTables t;
  float* out[MS];                 // downsampled tensors (may be null when only the loss terms are wanted)
  const float* pred[MS];          // predictions to compare with (null: plain downsample)
  float* dir[MS];                 // (pred - down) / epe per element (may be null)
  float* epe;                     // [ns] per-scale sum of end-point errors, divided by N (null without pred)
  const float* edge[MS];          // HFEM 'edges' loss: [N,oH,oW,1] edge map of scale s (null: plain EPE), and its
  float lambda_w;                 // weight: every pixel's error counts (1 + lambda_w * edge) times (utils.py:326-334)
  size_t ctrlOff, partOff;        // byte offsets in the workspace: control block (ticket + block sums), partials
};

// Shared-memory layout of the band kernel (bytes); used by the host for the launch size.
struct BandSmem {
  size_t band, rowW, rowD, fmask, bar, total;
};
__host__ __device__ inline size_t up16(size_t v) { return (v + 15) & ~(size_t)15; }
__host__ __device__ inline BandSmem band_smem_layout(int rowfloats) {
  BandSmem L;
  size_t o = 0;
  L.band = o; o = up16(o + (size_t)RB * rowfloats * 4);
  L.rowW = o; o = up16(o + (size_t)MS * RB * 8);
  L.rowD = o; o = up16(o + (size_t)MS * RB * 4);
  L.fmask = o; o = up16(o + (size_t)RB * 4);
  L.bar = o; o = up16(o + 8);
  L.total = o;
  return L;
}

__device__ __forceinline__ uint32_t ms_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

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
// CT channels; PXV = 4 / CT pixels per 16-byte vector; a thread owns at most NVEC vectors (P pixels) of every row.
// Partial entry of (band, scale, output row, interval dx): [A values x CT, A NaN weights x CT, B values, B NaN weights].
template <int CT, bool SCALED>
__global__ void __launch_bounds__(1024, 1) ms_band_kernel(MsParams p, const float* __restrict__ in,
                                                          float* __restrict__ ws) {
  constexpr int PXV = 4 / CT;
  constexpr int NVEC = CT == 2 ? 5 : 3;
  constexpr int P = NVEC * PXV;
  constexpr int NE = 4 * CT;
  extern __shared__ __align__(128) uint8_t sm[];
  const int tid = threadIdx.x, lane = tid & 31;
  const int band = blockIdx.x, n = blockIdx.y;
  const int yb = band * RB, ye = min(p.iH - 1, yb + RB - 1);
  const int nrows = ye - yb + 1;
  const int iW = p.iW, rowf = iW * CT;
  const int* wsi = reinterpret_cast<const int*>(ws);
  const BandSmem Ls = band_smem_layout(rowf);
  const float* bandbuf = reinterpret_cast<const float*>(sm + Ls.band);
  float2* rowW = reinterpret_cast<float2*>(sm + Ls.rowW);     // [MS][RB]
  int* rowD = reinterpret_cast<int*>(sm + Ls.rowD);           // [MS][RB]
  int* fmask = reinterpret_cast<int*>(sm + Ls.fmask);         // [RB] bit s: scale s finishes an output row before row i
  const uint32_t bar = ms_smem_u32(sm + Ls.bar);

  // ---- the band, by the bulk-copy engine: one request per 16 rows at most, one barrier ----
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const uint32_t bytes = (uint32_t)nrows * (uint32_t)rowf * 4u;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    const float* src = in + ((size_t)n * p.iH + yb) * rowf;
    uint32_t done = 0;
    while (done < bytes) {
      const uint32_t chunk = min(bytes - done, 32768u);
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(ms_smem_u32(sm + Ls.band) + done), "l"(reinterpret_cast<const char*>(src) + done), "r"(chunk),
                     "r"(bar) : "memory");
      done += chunk;
    }
  }

  // ---- which interval, which run ----
  int s = 0;
  while (s + 1 < p.ns && tid >= p.tOff[s + 1]) ++s;           // warp-uniform: tOff are multiples of 32
  const int local = tid - p.tOff[s];
  const int J = p.J[s], Lr = p.L[s];
  const int dx = local / J, j = local & (J - 1);
  const int oW = p.oW[s];
  const bool active = dx < oW;
  float wA[P], wB[P];
  int vec0 = 0, nvec = 0;
  {
    const int* cs = wsi + p.t.colS[s];
    const int c0 = active ? __ldg(cs + dx) : 0, c1 = active ? __ldg(cs + dx + 1) : 0;
    const int rr = (c0 & ~(PXV - 1)) + j * Lr;               // this lane's share of the interval: [rr, rr + Lr)
    const int lo = max(c0, rr), hi = min(c1, rr + Lr);
    const int r0 = lo & ~(PXV - 1);                           // the 16-byte vector that holds its first column
    const float2* cw = reinterpret_cast<const float2*>(ws + p.t.colW[s]);
#pragma unroll
    for (int k = 0; k < P; ++k) {
      const int px = r0 + k;
      const bool in_run = active && px >= lo && px < hi;
      const float2 w = in_run ? __ldg(cw + px) : make_float2(0.f, 0.f);
      wA[k] = w.x; wB[k] = w.y;
    }
    vec0 = r0 / PXV;
    nvec = (active && hi > lo) ? (hi - r0 + PXV - 1) / PXV : 0;   // <= Lr / PXV: rr is vector aligned
  }

  // ---- this band's slice of the row tables ----
  for (int idx = tid; idx < MS * RB; idx += blockDim.x) {
    const int sc = idx / RB, i = idx - sc * RB;
    float2 w = make_float2(0.f, 0.f);
    int d = 0;
    if (sc < p.ns && i < nrows) {
      w = __ldg(reinterpret_cast<const float2*>(ws + p.t.rowW[sc]) + yb + i);
      d = __ldg(wsi + p.t.rowD[sc] + yb + i);
    }
    rowW[idx] = w; rowD[idx] = d;
  }
  __syncthreads();                                            // also: the barrier is initialised before anyone polls it
  for (int i = tid; i < RB; i += blockDim.x) {
    int m = 0;
    if (i > 0 && i < nrows)
      for (int sc = 0; sc < p.ns; ++sc) m |= (rowD[sc * RB + i] != rowD[sc * RB + i - 1]) ? (1 << sc) : 0;
    fmask[i] = m;
  }
  __syncthreads();
  {
    uint32_t ok = 0;
    while (!ok) {
      asm volatile("{\n.reg .pred q;\nmbarrier.try_wait.parity.shared::cta.b64 q, [%1], 0;\nselp.u32 %0, 1, 0, q;\n}"
                   : "=r"(ok) : "r"(bar) : "memory");
    }
  }

  const float scale = p.in_scale;
  float* part = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + p.partOff) +
                (((size_t)n * p.nbands + band) * p.pTile + p.pOff[s]) * NE;
  const int d_first = rowD[s * RB];
  const float4* vbase = reinterpret_cast<const float4*>(bandbuf) + vec0;
  const int rowv = rowf / 4;                                  // 16-byte vectors per row

  // x-weighted sums of one row, careful path: NaN samples are left out of the values and counted in the NaN weights
  auto careful_row = [&](int i, float (&hA)[CT], float (&hB)[CT], float (&nA)[CT], float (&nB)[CT]) {
#pragma unroll
    for (int c = 0; c < CT; ++c) hA[c] = hB[c] = nA[c] = nB[c] = 0.f;
#pragma unroll
    for (int k = 0; k < NVEC; ++k) {
      if (k < nvec) {
        const float4 v = vbase[(size_t)i * rowv + k];
        const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int q = 0; q < PXV; ++q) {
          const float a = wA[k * PXV + q], b = wB[k * PXV + q];
          if (a == 0.f && b == 0.f) continue;                 // outside the run (or beyond every centre): not a sample
#pragma unroll
          for (int c = 0; c < CT; ++c) {
            const float x = SCALED ? __fmul_rn(e[q * CT + c], scale) : e[q * CT + c];
            if (x == x) { hA[c] = fmaf(a, x, hA[c]); hB[c] = fmaf(b, x, hB[c]); }
            else { nA[c] += a; nB[c] += b; }
          }
        }
      }
    }
  };

  float accA[2][CT], accB[2][CT];
#pragma unroll
  for (int k = 0; k < 2; ++k)
#pragma unroll
    for (int c = 0; c < CT; ++c) accA[k][c] = accB[k][c] = 0.f;
  unsigned marked = 0;                                        // bit i: row i of the band took the careful path here

  // stores output row dy (live row k of the accumulators) of this thread's interval
  auto flush = [&](int k, int dy) {
    if (dy >= p.oH[s]) return;                                // warp-uniform
    float v[NE];
#pragma unroll
    for (int c = 0; c < CT; ++c) { v[c] = accA[k][c]; v[CT + c] = 0.f; v[2 * CT + c] = accB[k][c]; v[3 * CT + c] = 0.f; }
    if (__any_sync(FULL, marked != 0)) {                      // rare: rebuild the NaN weights from the marked rows
      for (int i = 0; i < nrows; ++i) {
        if (!((marked >> i) & 1u)) continue;
        const int d = rowD[s * RB + i];
        const float2 w = rowW[s * RB + i];
        const float wy = d == dy ? w.x : (d + 1 == dy ? w.y : 0.f);
        if (wy == 0.f) continue;
        float hA[CT], hB[CT], nA[CT], nB[CT];
        careful_row(i, hA, hB, nA, nB);
#pragma unroll
        for (int c = 0; c < CT; ++c) { v[CT + c] = fmaf(wy, nA[c], v[CT + c]); v[3 * CT + c] = fmaf(wy, nB[c], v[3 * CT + c]); }
      }
    }
    for (int off = J >> 1; off >= 1; off >>= 1) {             // the J lanes of an interval, a fixed tree
#pragma unroll
      for (int e = 0; e < NE; ++e) v[e] += __shfl_xor_sync(FULL, v[e], off);
    }
    if (active && j == 0) {
      float* e = part + ((size_t)(dy - d_first) * oW + dx) * NE;
      *reinterpret_cast<float4*>(e) = make_float4(v[0], v[1], v[2], v[3]);
      if constexpr (NE == 8) *reinterpret_cast<float4*>(e + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
  };

  // ---- walk the rows ----
  for (int i = 0; i < nrows; ++i) {
    if ((fmask[i] >> s) & 1) {                                // warp-uniform: this scale passed the centre of its next row
      flush(0, rowD[s * RB + i - 1]);
#pragma unroll
      for (int c = 0; c < CT; ++c) { accA[0][c] = accA[1][c]; accA[1][c] = 0.f; accB[0][c] = accB[1][c]; accB[1][c] = 0.f; }
    }
    float hA[CT], hB[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) hA[c] = hB[c] = 0.f;
#pragma unroll
    for (int k = 0; k < NVEC; ++k) {
      if (k < nvec) {
        const float4 v = vbase[(size_t)i * rowv + k];
        // flow * in_scale is a rounded fp32 product, as if the scaled tensor had been materialised (flownet_c.py:128)
        const float e[4] = {SCALED ? __fmul_rn(v.x, scale) : v.x, SCALED ? __fmul_rn(v.y, scale) : v.y,
                            SCALED ? __fmul_rn(v.z, scale) : v.z, SCALED ? __fmul_rn(v.w, scale) : v.w};
#pragma unroll
        for (int q = 0; q < PXV; ++q)
#pragma unroll
          for (int c = 0; c < CT; ++c) {
            hA[c] = fmaf(wA[k * PXV + q], e[q * CT + c], hA[c]);
            hB[c] = fmaf(wB[k * PXV + q], e[q * CT + c], hB[c]);
          }
      }
    }
    bool bad = false;
#pragma unroll
    for (int c = 0; c < CT; ++c) bad |= !(hA[c] == hA[c]) || !(hB[c] == hB[c]);
    if (bad) {                                                // a NaN sample, or an infinity next to a zero weight
      float nA[CT], nB[CT];
      careful_row(i, hA, hB, nA, nB);
      marked |= 1u << i;
    }
    const float2 wy = rowW[s * RB + i];
#pragma unroll
    for (int c = 0; c < CT; ++c) {
      accA[0][c] = fmaf(wy.x, hA[c], accA[0][c]); accA[1][c] = fmaf(wy.y, hA[c], accA[1][c]);
      accB[0][c] = fmaf(wy.x, hB[c], accB[0][c]); accB[1][c] = fmaf(wy.y, hB[c], accB[1][c]);
    }
  }
  const int cur = rowD[s * RB + nrows - 1];
  flush(0, cur);
  flush(1, cur + 1);
}

// --------------------------------------------------------------------------------------------------- combine
template <int CT>
__global__ void __launch_bounds__(CB) ms_combine_kernel(MsParams p, float* __restrict__ ws) {
  constexpr int NV = 2 * CT;
  __shared__ float red[CB / 32];
  __shared__ bool lastBlock;
  __shared__ int dfirst[MAXBANDS];
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
  // first output row each band touches (the row its partial slots count from)
  for (int b = tid; b < p.nbands; b += CB) dfirst[b] = __ldg(wsi + p.t.rowD[s] + b * RB);
  __syncthreads();
  float epe = 0.f;
  if (live) {
    const int dy = idx / oW, dx = idx - dy * oW;
    const int2 br = __ldg(reinterpret_cast<const int2*>(wsi + p.t.bandR[s]) + dy);
    float acc[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) acc[k] = 0.f;
    const float* base = parts + ((size_t)n * p.nbands * p.pTile + p.pOff[s] + dx) * (2 * NV);
    const size_t bstride = (size_t)p.pTile * (2 * NV);
#pragma unroll 4
    for (int b = br.x; b <= br.y; ++b) {                     // the bands that feed this row, in band order
      const float* e = base + b * bstride + (size_t)(dy - dfirst[b]) * oW * (2 * NV);
      add_entry<NV>(acc, e);                                 // A part of this column's own interval
      if (dx > 0) add_entry<NV>(acc, e - NV);                // B part of the interval to its left
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
      // HFEM 'edges' (utils.py:326-334): epe + lambda * (epe * edges) per pixel, each product rounded as the graph's
      // separate nodes round it; the gradient w.r.t. the prediction carries the factor 1 + lambda * edges
      const float edge = p.edge[s] ? __ldg(p.edge[s] + ((size_t)n * oH + dy) * oW + dx) : 0.f;
      if (p.dir[s]) {
        const float w = fmaf(p.lambda_w, edge, 1.0f);
#pragma unroll
        for (int c = 0; c < CT; ++c)                          // 0/0 = NaN, as the gradient of sqrt at 0
          p.dir[s][o + c] = p.edge[s] ? diff[c] / epe * w : diff[c] / epe;
      }
      if (p.edge[s]) epe = __fadd_rn(epe, __fmul_rn(p.lambda_w, __fmul_rn(epe, edge)));
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
  if (ns < 1 || ns > MS || N > 65535) return false;
  const int P = C == 2 ? 10 : 12;                  // pixels per lane and row at most (NVEC 16-byte vectors)
  p.N = N; p.iH = iH; p.iW = iW; p.ns = ns;
  p.nbands = (iH + RB - 1) / RB;
  if (p.nbands > MAXBANDS) return false;
  int words = 0, pt = 0, blk = 0, thr = 0;
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
    // an interval holds at most floor(ws) + 1 columns; its first lane starts up to px - 1 columns before it
    const int span = (int)p.ws[s] + 2 + (px - 1);
    int J = 1, L = 0;
    for (;; J *= 2) {
      L = ((span + J - 1) / J + px - 1) / px * px;
      if (L <= P) break;
      if (J == 32) return false;
    }
    p.J[s] = J; p.L[s] = L;
    p.tOff[s] = thr;
    thr += (oW * J + 31) / 32 * 32;
    p.pOff[s] = pt;
    pt += p.maxRows[s] * oW;
    p.blkOff[s] = blk;
    blk += (oH * oW + CB - 1) / CB;
  }
  if (thr > 1024) return false;
  p.nthr = thr;
  for (int s = ns; s < MS; ++s) {   // unused scales
    p.oH[s] = p.oW[s] = 2; p.J[s] = 1; p.L[s] = px; p.maxRows[s] = 0; p.pOff[s] = pt; p.hs[s] = p.ws[s] = 2.f;
  }
  for (int s = ns; s <= MS; ++s) { p.blkOff[s] = blk; p.tOff[s] = thr; }
  p.pTile = pt;
  // table area
  for (int s = 0; s < ns; ++s) { p.t.rowW[s] = take(2 * iH); p.t.rowD[s] = take(iH); }
  for (int s = 0; s < ns; ++s) p.t.colW[s] = take(2 * iW);
  for (int s = 0; s < ns; ++s) p.t.colS[s] = take(p.oW[s] + 1);
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
  *ws_bytes = p.partOff + (size_t)N * p.nbands * p.pTile * (4 * C) * sizeof(float);
  *band_smem = band_smem_layout(iW * C).total;
  return *band_smem <= 100 * 1024;                 // the whole band stays resident (two CTAs per SM at least)
}

template <int CT>
int ms_launch(const MsParams& p, size_t band_smem, const float* in, void* ws, cudaStream_t st) {
  static PerDeviceOnce once;   // per template instantiation; function attributes are per device
  const cudaError_t attr_err = per_device_once(once, [] {
    cudaError_t e = cudaFuncSetAttribute(ms_band_kernel<CT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         100 * 1024);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(ms_band_kernel<CT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    return e;
  });
  if (attr_err != cudaSuccess) { note_cuda_error(attr_err); (void)cudaGetLastError(); return FN2_E_CUDA; }
  float* w = static_cast<float*>(ws);
  ms_tables_kernel<<<dim3(TBX, (unsigned)p.ns), 256, 0, st>>>(p, w);
  int rc = after_launch();
  if (rc != FN2_OK) return rc;
  const dim3 grid((unsigned)p.nbands, (unsigned)p.N);
  if (p.in_scale == 1.0f)
    ms_band_kernel<CT, false><<<grid, p.nthr, band_smem, st>>>(p, in, w);
  else
    ms_band_kernel<CT, true><<<grid, p.nthr, band_smem, st>>>(p, in, w);
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

// Whether EVERY group of (up to five) scales takes the fused kernels: what fn2_multiscale_epe_forward needs.
bool multiscale_fused_supported(const int* out_hw, int num_scales, int N, int iH, int iW, int C) {
  for (int s0 = 0; s0 < num_scales; s0 += MS) {
    const int ns = num_scales - s0 < MS ? num_scales - s0 : MS;
    MsParams p = {};
    size_t wsb = 0, smem = 0;
    if (!ms_plan(&p, out_hw + 2 * s0, ns, N, iH, iW, C, &wsb, &smem)) return false;
  }
  return true;
}

// predictions / residual_dir / epe_sums may be null (plain multi-scale downsample); outs entries may be null when
// predictions are given (only the loss terms wanted).
int multiscale_forward(const float* in, float in_scale, float* const* outs, const float* const* preds,
                       float* const* dirs, float* epe_sums, const int* out_hw, int num_scales, int N, int iH, int iW,
                       int C, void* ws, size_t ws_bytes, cudaStream_t st, const float* const* edges, float lambda_w) {
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
      p.edge[s] = (preds && edges) ? edges[s0 + s] : nullptr;
    }
    p.lambda_w = lambda_w;
    p.epe = preds ? epe_sums + s0 : nullptr;
    const int rc = C == 2 ? ms_launch<2>(p, smem, in, ws, st) : ms_launch<1>(p, smem, in, ws, st);
    if (rc != FN2_OK) return rc;
  }
  return FN2_OK;
}

// ------------------------------------------------------------------------------------------- built-in scratch
// fn2_downsample_multi_forward_noalloc (fn2ops.h): the same three kernels, for callers that cannot hand over a
// workspace because they have no way of getting device memory (a ctypes stub inside a TensorFlow process, whose
// allocator has taken the GPU's memory for itself and is not reachable from Python; any plug-in host that passes inputs
// and outputs only).  The scratch then lives in the library image: every GPU that loads libfn2ops.so carries these
// bytes until the process exits, whether the entry point is used or not.
//
// g_ms_builtin_scratch is a DEVICE array.  Kernels of this file could name it directly; host code cannot: on the host
// the identifier is only a handle for the runtime's symbol calls, and handing it to a kernel as if it were a pointer
// would make the kernel fault (or scribble over host memory where the GPU can reach it).  The address that kernels can
// use comes from cudaGetSymbolAddress, differs from device to device, and is looked up once per device ordinal by
// whichever call gets there first, with that device current.
namespace {

constexpr size_t MS_BUILTIN_BYTES = FN2_DOWNSAMPLE_MULTI_BUILTIN_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_ms_builtin_scratch[MS_BUILTIN_BYTES];

std::atomic<void*> g_ms_builtin_ptr[FN2_MAX_DEVICES];   // device address of the array, per device ordinal (null = not asked yet)

int ms_builtin_scratch(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { note_cuda_error(e); (void)cudaGetLastError(); return FN2_E_CUDA; }
  const bool tracked = dev >= 0 && dev < FN2_MAX_DEVICES;
  void* ptr = tracked ? g_ms_builtin_ptr[dev].load(std::memory_order_acquire) : nullptr;
  if (!ptr) {
    // not an allocation: the array is part of the module; this asks where the module put it on the current device.
    // Two threads racing here get the same answer.
    e = cudaGetSymbolAddress(&ptr, g_ms_builtin_scratch);
    if (e != cudaSuccess || !ptr) {
      note_cuda_error(e != cudaSuccess ? e : cudaErrorInvalidSymbol);
      (void)cudaGetLastError();
      return FN2_E_CUDA;
    }
    if (tracked) g_ms_builtin_ptr[dev].store(ptr, std::memory_order_release);
  }
  *out = ptr;
  return FN2_OK;
}

}  // namespace

// The array keeps the previous call's tables, partial sums and block sums.  Nothing of that is read before this call
// has rewritten it: ms_tables_kernel rebuilds every table entry and zeroes the combine ticket (the one word that must
// start at zero) as its first act, in stream order ahead of the band and combine kernels; the band kernel writes
// every partial the combine kernel reads.  That holds only while calls do not overlap on the GPU (see fn2ops.h).
int multiscale_forward_noalloc(const float* in, float* const* outs, const int* out_hw, int num_scales, int N, int iH,
                               int iW, int C, cudaStream_t st) {
  if (multiscale_workspace_bytes(out_hw, num_scales, N, iH, iW, C) > MS_BUILTIN_BYTES) return FN2_E_BUILTIN_SCRATCH;
  void* ws = nullptr;
  const int rc = ms_builtin_scratch(&ws);
  if (rc != FN2_OK) return rc;
  return multiscale_forward(in, 1.0f, outs, nullptr, nullptr, nullptr, out_hw, num_scales, N, iH, iW, C, ws,
                            MS_BUILTIN_BYTES, st);
}

}  // namespace fn2
