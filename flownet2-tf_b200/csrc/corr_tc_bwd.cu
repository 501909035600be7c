// corr_tc_bwd.cu -- FlowNetC correlation BACKWARD on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), sm_100a.
//
// Formulation (kernel_size 1, stride_1 1, pad == max_displacement; flownet_c.py:40, correlation_grad_kernel.cu.cc):
//   da[i,c] = 1/C * sum_j G[i,j] * b[j,c]          db[j,c] = 1/C * sum_i G[i,j] * a[i,c]
// where, inside one (y%s2, x%s2) parity class, i = (ya,xa) and j = (yb,xb) are pixels of the half-resolution grid and
//   G[i,j] = grad[n, i, (yb-ya+r)*gw + (xb-xa+r)]  if |yb-ya| <= r and |xb-xa| <= r, else 0
// is the banded "gradient Gram matrix".  Both gradients are dense GEMMs against G, so no scatter and no atomics:
// every output element is produced by exactly one CTA.
//
// One CTA owns a tile of 8 class-rows x 32 class-cols of OUTPUT pixels (a-pixels for da: mode 0, b-pixels for db:
// mode 1) and all C channels, and walks over the <= 8+2r source rows t of the other tensor X (b for da, a for db):
//   D^T[c, out pixel] += X_t^T[c, src pixel] * Gblk_t[src pixel, out pixel]       (M = 128 channels, N = 256, K = 32)
//   * X_t (32 pixels x C channels, NHWC) arrives by TMA as C/32 chunks of [32 pixels][32 channels] (swizzle
//     128B_ATOM_32B, the one layout the tensor core accepts for an MN-major 32-bit operand); read as the MN-major A
//     operand (channels contiguous) -- no transpose, no conversion: fp32 bits are consumed as tf32.
//   * Gblk_t is built by sixteen warps straight from `gradients` into shared memory in the same swizzled layout: eight
//     32x32 blocks (one per output row of the tile), rows = a-pixel, cols = b-pixel, each row the 21-float run
//     grad[i, p, :] shifted to its diagonal position.  The same bytes are a K-major B operand for da (N = a-pixel
//     rows, K = b-pixel cols) and an MN-major B operand for db (N = b-pixel cols, K = a-pixel rows).
//   * the accumulator D^T (2 halves of 128 channels x 256 output pixels, fp32) fills the 512 TMEM columns; after the
//     last source row the same sixteen warps read it back (lane = channel) and store 128-byte channel runs.
// CTAs of the two modes for the same image are adjacent in the grid so the second read of `gradients` hits L2.
//
// Wide images (class width > 32): output tiles are 32 columns wide and need source columns x0-10 .. x0+41, so each
// source row is visited as two 32-pixel items with origins x0-16 and x0+16 (TMA zero fill outside the image).
// Row halos (strip partitioning, fn2_correlation_backward_halo): b and db may be taller than a; a-row y meets b-rows
// y + boff + p in class space.  db is produced for every row of the tall tensor (halo rows = the partial sums the
// neighbouring strip adds to its own).
//
// Warp roles (576 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2-17 = G builders, then
// epilogue.
//
// fp32-grade mode (X3, FN2_PREC_TF32X3): both operands are split as x = hi + lo (hi = the 19 bits the tensor core
// reads from a raw fp32 word, lo = x - hi; fn2_tc.cuh x3_lo) and every step issues THREE MMAs: X.G into the main
// accumulator, X_lo.G + X.G_lo into a second one (the lo.lo term, ~2^-21 relative, is dropped).  The tensor core adds
// into its fp32 accumulator with truncation, so the error of a sum grows with the number of accumulation steps: the
// small terms get their own 256 TMEM columns (their rounding is 2^-11 times smaller) and the main accumulator takes
// a third of the steps; the epilogue adds the two in fp32 on the CUDA cores.  The builders write G_lo next to
// G from the value they already hold in a register, and convert the TMA-landed X chunk to X_lo in shared memory.  To
// make room a CTA takes only 128 of the channels (grid x ceil(C/128)): a stage is X 16 KB + X_lo 16 KB + G 32 KB +
// G_lo 32 KB, two stages, and each of the two accumulators needs 256 TMEM columns.  Results meet the fp32 tolerance
// (1e-5 relative + 1e-6 absolute against a float64 evaluation).
#include <type_traits>
#include "fn2_tc.cuh"

namespace fn2 {
namespace {

using namespace tc;

constexpr int TILE_R = 8;                 // output rows per tile (class space)
constexpr int TILE_C = 32;                // output cols per tile
constexpr int XCHUNK = 32 * 128;          // one X chunk: 32 pixels x 32 channels fp32 = 4 KB
constexpr int MAX_CH = 8;                 // C <= 256
constexpr int XSTAGE = MAX_CH * XCHUNK;   // 32 KB per stage
constexpr int GBLOCK = 32 * 128;          // one G block: 32 a-pixels x 32 b-pixels fp32 = 4 KB
constexpr int GSTAGE = TILE_R * GBLOCK;   // 32 KB per stage
constexpr int NBUILD = 2 * TILE_R;        // builder / epilogue warps: two per output row of the tile
constexpr int NTHREADS = 64 + 32 * NBUILD;
constexpr int TILE_N = TILE_R * TILE_C;    // output pixels per tile = N of the MMA (256)
constexpr int TMEM_COLS = 2 * TILE_N;      // two 128-channel halves of the accumulator

struct BwdParams {
  CUtensorMap x[2][4];  // [mode][class]: mode 0 (da) streams b, mode 1 (db) streams a
  CUtensorMap o[2][4];  // [mode][class]: the output tensor (da, db), boxes of 32 channels x 32 class pixels, no swizzle
  const float* grad;
  const float* mask;    // FUSED: stored activations, indexed like grad (LeakyReLU derivative = sign test); else unused
  int gpitch;           // floats between consecutive pixels of grad / mask (D for a plain gradient tensor)
  float slope;
  float* out[2];        // da, db
  int N, H, W, C, s2, ncls;
  int Hc[2], Wc[2];
  int Hcb[2];           // class-space rows of b / db (>= Hc + boff with row halos)
  int bH;               // full-resolution rows of b / db
  int boff;             // class-space row of b that a-row 0 meets at displacement 0
  int TY, TX;
  int nchunk;           // C / 32
  int nhalf;            // ceil(C / 128)
  float inv;            // 1 / C
};

template <int NST>
struct __align__(8) BwdBarriers {
  uint64_t x_full[NST];
  uint64_t g_full[NST];
  uint64_t empty[NST];
  uint64_t acc_full;
  uint32_t tmem_base;
};

template <int NST, bool X3>
constexpr size_t bwd_smem_bytes() {
  return 1024 + (size_t)NST * (XSTAGE + (X3 ? 2 : 1) * GSTAGE) + sizeof(BwdBarriers<NST>);
}

template <int NST, int S2, int R, bool X3, bool FUSED>
__global__ void __launch_bounds__(NTHREADS, 1) corr_tc_bwd_kernel(const __grid_constant__ BwdParams P) {
  static_assert(NST >= 2 && NST <= 4, "the G builders unroll the stage ring explicitly");
  constexpr int GW = 2 * R + 1;
  constexpr int D = GW * GW;
  // X3: the X stage holds 4 raw chunks (this CTA's 128 channels) then their 4 lo chunks -- the same 32 KB, addressed
  // like the two channel halves of the plain kernel; the G stage holds G then G_lo.
  constexpr int GST = X3 ? 2 * GSTAGE : GSTAGE;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sX = smem;
  uint8_t* sG = sX + (size_t)NST * XSTAGE;
  BwdBarriers<NST>* bars = reinterpret_cast<BwdBarriers<NST>*>(sG + (size_t)NST * GST);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- which tile: ((((n * ncls + cls) * TY + ty) * TX + tx) * nhalf + ch) * 2 + mode   (ch: X3 only) ----
  int t = blockIdx.x;
  const int mode = t & 1; t >>= 1;
  int ch = 0;                                     // X3: which 128 channels this CTA produces
  if (X3) { ch = t % P.nhalf; t /= P.nhalf; }
  const int c_first = ch * 4;                     // first 32-channel chunk
  const int nch_x = X3 ? min(4, P.nchunk - c_first) : P.nchunk;   // chunks this CTA streams
  const int tx = t % P.TX; t /= P.TX;
  const int ty = t % P.TY; t /= P.TY;
  const int cls = t % P.ncls;
  const int n = t / P.ncls;
  const int cy = cls / P.s2, cx = cls % P.s2;
  const int Wc = P.Wc[cx];
  const int Hout = mode ? P.Hcb[cy] : P.Hc[cy];   // class rows of the output tensor (da: a-space, db: b-space)
  const int Hsrc = mode ? P.Hc[cy] : P.Hcb[cy];   // class rows of the streamed tensor
  const int Ha = P.Hc[cy];
  const int r0 = ty * TILE_R, x0 = tx * TILE_C;
  if (r0 >= Hout || x0 >= Wc) return;  // whole CTA leaves together, before any barrier exists
  // output row ro meets source rows ro + shift - R .. ro + shift + R  (shift = +boff for da, -boff for db)
  const int shift = mode ? -P.boff : P.boff;
  const int t_first = max(0, r0 + shift - R);
  const int t_last = min(Hsrc - 1, r0 + TILE_R - 1 + shift + R);
  const int nrows = t_last - t_first + 1;
  const int nchunk = P.nchunk;
  const int nhx = P.TX > 1 ? 2 : 1;               // 32-pixel source items per source row
  const int nitems = nrows * nhx;
  const int ncol = min(TILE_C, Wc - x0);
  float* const outp = P.out[mode];
  const int Hfull = mode ? P.bH : P.H;            // full-resolution rows of the output tensor
  if (nrows <= 0) {  // no source row reaches this tile (halo deeper than the displacement range): zeros
    for (int qq = 0; qq < TILE_R; ++qq) {
      const int row = r0 + qq;
      if (row >= Hout) break;
      float* orow = outp + (((size_t)n * Hfull + (size_t)(row * P.s2 + cy)) * P.W + cx + (size_t)x0 * P.s2) * P.C;
      if (X3) {   // this CTA's channels only
        const int cw = nch_x * KCH;
        for (int e = threadIdx.x; e < ncol * cw; e += NTHREADS)
          orow[(size_t)(e / cw) * P.s2 * P.C + c_first * KCH + e % cw] = 0.f;
      } else {
        for (int e = threadIdx.x; e < ncol * P.C; e += NTHREADS) orow[(size_t)(e / P.C) * P.s2 * P.C + e % P.C] = 0.f;
      }
    }
    return;
  }

  if (threadIdx.x == 0) {
    for (int i = 0; i < NST; ++i) {
      mbar_init(smem_u32(&bars->x_full[i]), 1);
      mbar_init(smem_u32(&bars->g_full[i]), NBUILD);
      mbar_init(smem_u32(&bars->empty[i]), 1);
    }
    mbar_init(smem_u32(&bars->acc_full), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(smem_u32(&bars->tmem_base), TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ================================ TMA producer: one source row of X per stage ================================
    if (elect_one()) {
      const CUtensorMap* map = &P.x[mode][cls];
      for (int it = 0; it < nitems; ++it) {
        const int stage = it % NST;
        const int row = t_first + it / nhx;
        const int xs0 = nhx > 1 ? x0 - 16 + 32 * (it % nhx) : 0;
        mbar_wait(smem_u32(&bars->empty[stage]), ((it / NST) & 1) ^ 1);
        const uint32_t full = smem_u32(&bars->x_full[stage]);
        mbar_expect_tx(full, (uint32_t)nch_x * XCHUNK);
        for (int c = 0; c < nch_x; ++c)
          tma_load_4d(smem_u32(sX + (size_t)stage * XSTAGE + (size_t)c * XCHUNK), map, full, (c_first + c) * KCH, xs0,
                      row, n);
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    if (elect_one()) {
      // A = X^T: MN-major (channels contiguous), M = 128 channels = 4 chunks (LBO = chunk stride), K = 8 pixel rows.
      // B = G  : mode 0 K-major (N = a-pixel rows, K = b-pixel cols); mode 1 MN-major (N = b-pixel cols, LBO = block
      //          stride, K = a-pixel rows).
      const uint32_t idesc = idesc_tf32_major(128, TILE_N, 1, mode);
      for (int it = 0; it < nitems; ++it) {
        const int stage = it % NST;
        const uint32_t par = (it / NST) & 1;
        mbar_wait(smem_u32(&bars->x_full[stage]), par);
        mbar_wait(smem_u32(&bars->g_full[stage]), par);
        tc_fence_after();
        const uint32_t x_addr = smem_u32(sX + (size_t)stage * XSTAGE);
        const uint32_t g_addr = smem_u32(sG + (size_t)stage * GST);
        if constexpr (X3) {
#pragma unroll
          for (int k8 = 0; k8 < 4; ++k8) {
            const uint64_t a_hi = desc_sw128b32_mnmajor(x_addr + k8 * 1024, XCHUNK, 512);
            const uint64_t a_lo = desc_sw128b32_mnmajor(x_addr + 4 * XCHUNK + k8 * 1024, XCHUNK, 512);
            const uint64_t b_hi = mode ? desc_sw128b32_mnmajor(g_addr + k8 * 1024, GBLOCK, 512)
                                       : desc_sw128_kmajor(g_addr + k8 * 32);
            const uint64_t b_lo = mode ? desc_sw128b32_mnmajor(g_addr + GSTAGE + k8 * 1024, GBLOCK, 512)
                                       : desc_sw128_kmajor(g_addr + GSTAGE + k8 * 32);
            umma_tf32(tmem_base, a_hi, b_hi, idesc, (uint32_t)((it | k8) != 0));             // main accumulator
            umma_tf32(tmem_base + TILE_N, a_lo, b_hi, idesc, (uint32_t)((it | k8) != 0));    // small terms
            umma_tf32(tmem_base + TILE_N, a_hi, b_lo, idesc, 1u);
          }
        } else {
          for (int h = 0; h < P.nhalf; ++h) {
#pragma unroll
            for (int k8 = 0; k8 < 4; ++k8) {
              const uint64_t adesc = desc_sw128b32_mnmajor(x_addr + h * 4 * XCHUNK + k8 * 1024, XCHUNK, 512);
              const uint64_t bdesc = mode ? desc_sw128b32_mnmajor(g_addr + k8 * 1024, GBLOCK, 512)
                                          : desc_sw128_kmajor(g_addr + k8 * 32);
              umma_tf32(tmem_base + (uint32_t)h * TILE_N, adesc, bdesc, idesc, (uint32_t)((it | k8) != 0));
            }
          }
        }
        umma_commit(smem_u32(&bars->empty[stage]));  // frees X and G of this stage once the MMAs have read them
      }
      umma_commit(smem_u32(&bars->acc_full));
    }
  } else {
    // ================================ G builders (warps 2..9), then epilogue ================================
    const int bw = warp - 2;
    const int q = bw >> 1;            // block = output row r0 + q of the tile
    const int xh = (bw & 1) * 16;     // this warp fills a-pixel rows xh .. xh+15 of the block
    const int xb = lane;              // column of the block = b-pixel
    // Everything that does not change from item to item is hoisted: the swizzled shared-memory address of each of
    // the thread's 16 elements, and (per 32-column source half hx) the validity mask of the 16 (a-pixel, b-pixel)
    // pairs.  The 16 global addresses of an item are base + rr * (S2*D - 1) floats (next a-pixel, displacement
    // one less), so they are immediates off one 64-bit base pointer.
    // (FUSED: the pixel pitch is a run-time value, the gradient being a channel window of the concat buffer)
    const int GP = FUSED ? P.gpitch : D;
    const int ESTRIDE = S2 * GP - 1;
    uint32_t saddr[16];
    uint32_t okmask[2] = {0u, 0u};
    const uint32_t sG_q = smem_u32(sG) + (uint32_t)q * GBLOCK;
#pragma unroll
    for (int rr = 0; rr < 16; ++rr) {
      const int x = xh + rr;
      // mode 0: K-major operand, SWIZZLE_128B (16-byte chunk ^ (row & 7)); mode 1: MN-major operand,
      // SWIZZLE_128B_BASE32B (32-byte unit ^ (row & 3))
      const int off = mode ? ((((xb >> 3) ^ (x & 3)) << 5) | ((xb & 7) << 2))
                           : ((((xb >> 2) ^ (x & 7)) << 4) | ((xb & 3) << 2));
      saddr[rr] = sG_q + (uint32_t)(x * 128 + off);
#pragma unroll
      for (int hx = 0; hx < 2; ++hx) {
        const int xs0 = nhx > 1 ? x0 - 16 + 32 * hx : 0;
        const int aabs = (mode ? xs0 : x0) + x, babs = (mode ? x0 : xs0) + xb;
        const int o = babs - aabs + R;
        const bool ok = (aabs >= 0) && (aabs < Wc) && (babs >= 0) && (babs < Wc) && (o >= 0) && (o < GW);
        okmask[hx] |= (ok ? 1u : 0u) << rr;
      }
    }
    const bool qrow_ok = (r0 + q < Hout);
    const float* gimg = P.grad + (size_t)n * P.H * P.W * GP;
    const ptrdiff_t moff = FUSED ? (P.mask - P.grad) : 0;   // same indexing, other tensor

    // The gather of one item (16 floats per thread, 84-byte runs of `gradients`) is a full L2/DRAM round trip, so
    // items it+1 .. it+NST-1 are fetched into registers before item it is waited for and stored.
    auto load_item = [&](int it, float (&v)[16]) {
      const int tt = t_first + it / nhx;
      const int hx = it % nhx;
      const int xs0 = nhx > 1 ? x0 - 16 + 32 * hx : 0;
      // a-row and displacement row of this block for this source row: b-row = a-row + boff + p
      const int ya = mode ? tt : r0 + q;
      const int p = mode ? (r0 + q - P.boff - tt) : (tt - P.boff - (r0 + q));
      const bool blk_ok = qrow_ok && (ya < Ha) && (p >= -R) && (p <= R);
      const uint32_t m = blk_ok ? okmask[hx] : 0u;
      // block rows are a-pixels, block columns are b-pixels; the tile owns the a-columns for da, the b-columns for db
      const int a0 = (mode ? xs0 : x0) + xh, b0 = (mode ? x0 : xs0) + xb;
      const float* base = gimg + ((ptrdiff_t)((blk_ok ? ya : 0) * S2 + cy) * P.W + cx) * GP + (blk_ok ? p + R : 0) * GW +
                          (ptrdiff_t)a0 * (S2 * GP) + (b0 - a0 + R);
#pragma unroll
      for (int rr = 0; rr < 16; ++rr) v[rr] = ((m >> rr) & 1u) ? __ldg(base + rr * ESTRIDE) : 0.f;
      if constexpr (FUSED) {   // times the derivative of the LeakyReLU fused into the forward (flownet_c.py:41)
#pragma unroll
        for (int rr = 0; rr < 16; ++rr)
          if ((m >> rr) & 1u) v[rr] = __ldg(base + moff + rr * ESTRIDE) > 0.f ? v[rr] : v[rr] * P.slope;
      }
    };
    {
      float vq[NST][16];   // register ring aligned with the stage ring: item it lives in vq[it % NST]
      load_item(0, vq[0]);
#pragma unroll
      for (int i = 1; i < NST - 1; ++i)
        if (i < nitems) load_item(i, vq[i]);
      for (int it0 = 0; it0 < nitems; it0 += NST) {
        const uint32_t par = ((it0 / NST) & 1) ^ 1;
        auto stage_body = [&](auto ST) {     // stage == ST: compile-time offsets into the G ring
          constexpr int st = decltype(ST)::value;
          const int it = it0 + st;
          if (it < nitems) {
            float (&cur)[16] = vq[st];
            if (it + NST - 1 < nitems) load_item(it + NST - 1, vq[(st + NST - 1) % NST]);
            mbar_wait(smem_u32(&bars->empty[st]), par);
#pragma unroll
            for (int rr = 0; rr < 16; ++rr)
              asm volatile("st.shared.f32 [%0+%2], %1;" ::"r"(saddr[rr]), "f"(cur[rr]), "n"(st * GST) : "memory");
            if constexpr (X3) {
#pragma unroll
              for (int rr = 0; rr < 16; ++rr)
                asm volatile("st.shared.f32 [%0+%2], %1;" ::"r"(saddr[rr]), "f"(x3_lo(cur[rr])), "n"(st * GST + GSTAGE)
                             : "memory");
              // X_lo from the raw X chunks TMA has landed for this item: 16 KB, two float4 per builder thread
              mbar_wait(smem_u32(&bars->x_full[st]), par ^ 1);
              uint8_t* xs = sX + (size_t)st * XSTAGE;
              const int o0 = (int)(threadIdx.x - 64) * 16;
              const float4 x0v = *reinterpret_cast<const float4*>(xs + o0);
              const float4 x1v = *reinterpret_cast<const float4*>(xs + o0 + NBUILD * 32 * 16);
              *reinterpret_cast<float4*>(xs + 4 * XCHUNK + o0) = x3_lo4(x0v);
              *reinterpret_cast<float4*>(xs + 4 * XCHUNK + o0 + NBUILD * 32 * 16) = x3_lo4(x1v);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the MMA
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&bars->g_full[st]));
          }
        };
        stage_body(std::integral_constant<int, 0>{});
        stage_body(std::integral_constant<int, 1>{});
        if constexpr (NST >= 3) stage_body(std::integral_constant<int, 2>{});
        if constexpr (NST == 4) stage_body(std::integral_constant<int, 3>{});
      }
    }

    // ---- epilogue: D^T[channel lane, output pixel column] -> out[pixel, channel] ----
    mbar_wait(smem_u32(&bars->acc_full), 0);
    tc_fence_after();
    const int qd = warp & 3;        // TMEM lane quadrant this warp may read
    const int e = bw >> 2;          // the NBUILD/4 warps of a quadrant take alternate output rows
    const float inv = P.inv;
    // One TMEM load is [32 channels (lanes)] x [32 output pixels of one row]: exactly one TMA store box.  The warp
    // writes it to shared memory as [pixel][channel] (lane = channel: conflict-free 128-byte rows) and one lane hands it
    // to the TMA unit, which also clips the columns and keeps the LSU out of the store path.  Staging: two 4 KB boxes
    // per warp in the drained X ring.
    const uint32_t stg = smem_u32(sX) + (uint32_t)bw * 8192u;
    const CUtensorMap* omap = &P.o[mode][cls];
    int nb = 0;
    for (int h = 0; h < (X3 ? 1 : P.nhalf); ++h) {
      const int c0 = (X3 ? ch : h) * 128 + qd * 32;
      if (c0 >= P.C) break;         // warp-uniform
      for (int qq = e; qq < TILE_R; qq += NBUILD / 4) {
        const int row = r0 + qq;
        if (row >= Hout) continue;  // warp-uniform
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(h * TILE_N + qq * 32), v);
        if constexpr (X3) {   // + the accumulator of the small terms, in two pieces (fewer live registers)
          float w[16];
          tmem_ld16(tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(TILE_N + qq * 32), w);
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] += w[j];
          tmem_ld16(tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(TILE_N + qq * 32 + 16), w);
#pragma unroll
          for (int j = 0; j < 16; ++j) v[16 + j] += w[j];
        }
        const uint32_t buf = stg + (uint32_t)(nb & 1) * 4096u;
        if (nb >= 2) {
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < 32; ++j)
          asm volatile("st.shared.f32 [%0], %1;" ::"r"(buf + (uint32_t)(lane * 4 + j * 128)), "f"(v[j] * inv)
                       : "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
          tma_store_4d(omap, buf, c0, x0, row, n);
          tma_store_commit();
        }
        ++nb;
      }
    }
    if (lane == 0) tma_store_wait_read<0>();   // shared memory stays valid until it was read
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ---- CTA-pair version of the fp32-grade mode (cta_group::2), C = 256 -------------------------------------------------
// In the kernel above the two CTAs that produce the two 128-channel halves of a tile each gather, split and store the
// whole gradient Gram matrix G (eight 32x32 blocks + their lo parts per source row): the builders' work per MMA is
// twice that of the plain tf32 kernel and only two 96 KB stages fit.  Here the two CTAs are one cluster and ONE
// tcgen05.mma.cta_group::2 (M = 256 channels, N = 256 output pixels) serves both: each CTA streams and splits the X
// chunks of ITS 128 channels and builds only FOUR of the eight G blocks (the N rows it supplies); a stage is 64 KB
// (X 16 + X_lo 16 + G 16 + G_lo 16) and three of them fit, so gathers run two items ahead.  g_full lives in the
// leader and collects the builder warps of both CTAs (X landed and split, G built); empty and acc_full are multicast
// commits.  Same results as the single-CTA kernel bit for bit (the same products accumulate in the same order).
constexpr int PAIR_NST = 3;
constexpr size_t bwd_pair_smem_bytes() { return 1024 + (size_t)PAIR_NST * (XSTAGE + GSTAGE) + sizeof(BwdBarriers<PAIR_NST>); }

template <int S2, int R, bool FUSED>
__global__ void __launch_bounds__(NTHREADS, 1) corr_tc_bwd_x3_pair_kernel(const __grid_constant__ BwdParams P) {
  constexpr int NST = PAIR_NST;
  constexpr bool X3 = true;
  constexpr int GW = 2 * R + 1;
  constexpr int D = GW * GW;
  // X3: the X stage holds 4 raw chunks (this CTA's 128 channels) then their 4 lo chunks -- the same 32 KB, addressed
  // like the two channel halves of the plain kernel; the G stage holds G then G_lo.
  constexpr int GHALF = GSTAGE / 2;            // this CTA's four G blocks
  constexpr int GST = 2 * GHALF;               // ... and their lo parts
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sX = smem;
  uint8_t* sG = sX + (size_t)NST * XSTAGE;
  BwdBarriers<NST>* bars = reinterpret_cast<BwdBarriers<NST>*>(sG + (size_t)NST * GST);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- which tile: ((((n * ncls + cls) * TY + ty) * TX + tx) * nhalf + ch) * 2 + mode   (ch: X3 only) ----
  const uint32_t rank = cluster_ctarank();       // = blockIdx.x & 1: which 128 channels
  int t = blockIdx.x >> 1;
  const int mode = t & 1; t >>= 1;
  const int ch = (int)rank;
  const int c_first = ch * 4;                     // first 32-channel chunk
  const int nch_x = min(4, P.nchunk - c_first);   // chunks this CTA streams
  const int tx = t % P.TX; t /= P.TX;
  const int ty = t % P.TY; t /= P.TY;
  const int cls = t % P.ncls;
  const int n = t / P.ncls;
  const int cy = cls / P.s2, cx = cls % P.s2;
  const int Wc = P.Wc[cx];
  const int Hout = mode ? P.Hcb[cy] : P.Hc[cy];   // class rows of the output tensor (da: a-space, db: b-space)
  const int Hsrc = mode ? P.Hc[cy] : P.Hcb[cy];   // class rows of the streamed tensor
  const int Ha = P.Hc[cy];
  const int r0 = ty * TILE_R, x0 = tx * TILE_C;
  if (r0 >= Hout || x0 >= Wc) return;  // whole CTA leaves together, before any barrier exists
  // output row ro meets source rows ro + shift - R .. ro + shift + R  (shift = +boff for da, -boff for db)
  const int shift = mode ? -P.boff : P.boff;
  const int t_first = max(0, r0 + shift - R);
  const int t_last = min(Hsrc - 1, r0 + TILE_R - 1 + shift + R);
  const int nrows = t_last - t_first + 1;
  const int nchunk = P.nchunk;
  const int nhx = P.TX > 1 ? 2 : 1;               // 32-pixel source items per source row
  const int nitems = nrows * nhx;
  const int ncol = min(TILE_C, Wc - x0);
  float* const outp = P.out[mode];
  const int Hfull = mode ? P.bH : P.H;            // full-resolution rows of the output tensor
  if (nrows <= 0) {  // no source row reaches this tile (halo deeper than the displacement range): zeros
    for (int qq = 0; qq < TILE_R; ++qq) {
      const int row = r0 + qq;
      if (row >= Hout) break;
      float* orow = outp + (((size_t)n * Hfull + (size_t)(row * P.s2 + cy)) * P.W + cx + (size_t)x0 * P.s2) * P.C;
      if (X3) {   // this CTA's channels only
        const int cw = nch_x * KCH;
        for (int e = threadIdx.x; e < ncol * cw; e += NTHREADS)
          orow[(size_t)(e / cw) * P.s2 * P.C + c_first * KCH + e % cw] = 0.f;
      } else {
        for (int e = threadIdx.x; e < ncol * P.C; e += NTHREADS) orow[(size_t)(e / P.C) * P.s2 * P.C + e % P.C] = 0.f;
      }
    }
    return;
  }

  if (threadIdx.x == 0) {
    for (int i = 0; i < NST; ++i) {
      mbar_init(smem_u32(&bars->x_full[i]), 1);
      mbar_init(smem_u32(&bars->g_full[i]), 2 * NBUILD);   // leader's: the builder warps of BOTH CTAs
      mbar_init(smem_u32(&bars->empty[i]), 1);
    }
    mbar_init(smem_u32(&bars->acc_full), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) tmem_alloc_pair(smem_u32(&bars->tmem_base), TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();          // the peer's barriers exist before anything is signalled on them
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ================================ TMA producer: one source row of X per stage ================================
    if (elect_one()) {
      const CUtensorMap* map = &P.x[mode][cls];
      for (int it = 0; it < nitems; ++it) {
        const int stage = it % NST;
        const int row = t_first + it / nhx;
        const int xs0 = nhx > 1 ? x0 - 16 + 32 * (it % nhx) : 0;
        mbar_wait(smem_u32(&bars->empty[stage]), ((it / NST) & 1) ^ 1);
        const uint32_t full = smem_u32(&bars->x_full[stage]);
        mbar_expect_tx(full, (uint32_t)nch_x * XCHUNK);
        for (int c = 0; c < nch_x; ++c)
          tma_load_4d(smem_u32(sX + (size_t)stage * XSTAGE + (size_t)c * XCHUNK), map, full, (c_first + c) * KCH, xs0,
                      row, n);
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (leader only) ================================
    if (rank == 0 && elect_one()) {
      // A = X^T: MN-major, M = 256 channels = 128 per CTA (4 chunks, LBO = chunk stride), K = 8 pixel rows.
      // B = G  : N = 256 output pixels = 128 per CTA (4 blocks); mode 0 K-major, mode 1 MN-major (LBO = block stride).
      const uint32_t idesc = idesc_tf32_major(256, TILE_N, 1, mode);
      for (int it = 0; it < nitems; ++it) {
        const int stage = it % NST;
        const uint32_t par = (it / NST) & 1;
        mbar_wait(smem_u32(&bars->g_full[stage]), par);   // both CTAs: X landed and split, G and G_lo built
        tc_fence_after();
        const uint32_t x_addr = smem_u32(sX + (size_t)stage * XSTAGE);
        const uint32_t g_addr = smem_u32(sG + (size_t)stage * GST);
#pragma unroll
        for (int k8 = 0; k8 < 4; ++k8) {
          const uint64_t a_hi = desc_sw128b32_mnmajor(x_addr + k8 * 1024, XCHUNK, 512);
          const uint64_t a_lo = desc_sw128b32_mnmajor(x_addr + 4 * XCHUNK + k8 * 1024, XCHUNK, 512);
          const uint64_t b_hi = mode ? desc_sw128b32_mnmajor(g_addr + k8 * 1024, GBLOCK, 512)
                                     : desc_sw128_kmajor(g_addr + k8 * 32);
          const uint64_t b_lo = mode ? desc_sw128b32_mnmajor(g_addr + GHALF + k8 * 1024, GBLOCK, 512)
                                     : desc_sw128_kmajor(g_addr + GHALF + k8 * 32);
          umma_tf32_pair(tmem_base, a_hi, b_hi, idesc, (uint32_t)((it | k8) != 0));             // main accumulator
          umma_tf32_pair(tmem_base + TILE_N, a_lo, b_hi, idesc, (uint32_t)((it | k8) != 0));    // small terms
          umma_tf32_pair(tmem_base + TILE_N, a_hi, b_lo, idesc, 1u);
        }
        umma_commit_pair(smem_u32(&bars->empty[stage]), 3);  // frees X and G of this stage in both CTAs
      }
      umma_commit_pair(smem_u32(&bars->acc_full), 3);
    }
  } else {
    // ================================ G builders (warps 2..9), then epilogue ================================
    const int bw = warp - 2;
    const int ql = bw >> 2;           // which of this CTA's four G blocks
    const int q = 4 * (int)rank + ql; // block = output row r0 + q of the tile
    const int xh = (bw & 3) * 8;      // this warp fills a-pixel rows xh .. xh+7 of the block
    const int xb = lane;              // column of the block = b-pixel
    // Everything that does not change from item to item is hoisted: the swizzled shared-memory address of each of
    // the thread's 16 elements, and (per 32-column source half hx) the validity mask of the 16 (a-pixel, b-pixel)
    // pairs.  The 16 global addresses of an item are base + rr * (S2*D - 1) floats (next a-pixel, displacement
    // one less), so they are immediates off one 64-bit base pointer.
    // (FUSED: the pixel pitch is a run-time value, the gradient being a channel window of the concat buffer)
    const int GP = FUSED ? P.gpitch : D;
    const int ESTRIDE = S2 * GP - 1;
    constexpr int NR = 8;             // elements per thread and item
    uint32_t saddr[NR];
    uint32_t okmask[2] = {0u, 0u};
    const uint32_t sG_q = smem_u32(sG) + (uint32_t)ql * GBLOCK;
#pragma unroll
    for (int rr = 0; rr < NR; ++rr) {
      const int x = xh + rr;
      // mode 0: K-major operand, SWIZZLE_128B (16-byte chunk ^ (row & 7)); mode 1: MN-major operand,
      // SWIZZLE_128B_BASE32B (32-byte unit ^ (row & 3))
      const int off = mode ? ((((xb >> 3) ^ (x & 3)) << 5) | ((xb & 7) << 2))
                           : ((((xb >> 2) ^ (x & 7)) << 4) | ((xb & 3) << 2));
      saddr[rr] = sG_q + (uint32_t)(x * 128 + off);
#pragma unroll
      for (int hx = 0; hx < 2; ++hx) {
        const int xs0 = nhx > 1 ? x0 - 16 + 32 * hx : 0;
        const int aabs = (mode ? xs0 : x0) + x, babs = (mode ? x0 : xs0) + xb;
        const int o = babs - aabs + R;
        const bool ok = (aabs >= 0) && (aabs < Wc) && (babs >= 0) && (babs < Wc) && (o >= 0) && (o < GW);
        okmask[hx] |= (ok ? 1u : 0u) << rr;
      }
    }
    const bool qrow_ok = (r0 + q < Hout);
    const uint32_t g_full0 = mapa_u32(smem_u32(&bars->g_full[0]), 0);
    const float* gimg = P.grad + (size_t)n * P.H * P.W * GP;
    const ptrdiff_t moff = FUSED ? (P.mask - P.grad) : 0;   // same indexing, other tensor

    // The gather of one item (16 floats per thread, 84-byte runs of `gradients`) is a full L2/DRAM round trip, so
    // items it+1 .. it+NST-1 are fetched into registers before item it is waited for and stored.
    auto load_item = [&](int it, float (&v)[NR]) {
      const int tt = t_first + it / nhx;
      const int hx = it % nhx;
      const int xs0 = nhx > 1 ? x0 - 16 + 32 * hx : 0;
      // a-row and displacement row of this block for this source row: b-row = a-row + boff + p
      const int ya = mode ? tt : r0 + q;
      const int p = mode ? (r0 + q - P.boff - tt) : (tt - P.boff - (r0 + q));
      const bool blk_ok = qrow_ok && (ya < Ha) && (p >= -R) && (p <= R);
      const uint32_t m = blk_ok ? okmask[hx] : 0u;
      // block rows are a-pixels, block columns are b-pixels; the tile owns the a-columns for da, the b-columns for db
      const int a0 = (mode ? xs0 : x0) + xh, b0 = (mode ? x0 : xs0) + xb;
      const float* base = gimg + ((ptrdiff_t)((blk_ok ? ya : 0) * S2 + cy) * P.W + cx) * GP + (blk_ok ? p + R : 0) * GW +
                          (ptrdiff_t)a0 * (S2 * GP) + (b0 - a0 + R);
#pragma unroll
      for (int rr = 0; rr < NR; ++rr) v[rr] = ((m >> rr) & 1u) ? __ldg(base + rr * ESTRIDE) : 0.f;
      if constexpr (FUSED) {   // times the derivative of the LeakyReLU fused into the forward (flownet_c.py:41)
#pragma unroll
        for (int rr = 0; rr < NR; ++rr)
          if ((m >> rr) & 1u) v[rr] = __ldg(base + moff + rr * ESTRIDE) > 0.f ? v[rr] : v[rr] * P.slope;
      }
    };
    {
      float vq[NST][NR];   // register ring aligned with the stage ring: item it lives in vq[it % NST]
      load_item(0, vq[0]);
#pragma unroll
      for (int i = 1; i < NST - 1; ++i)
        if (i < nitems) load_item(i, vq[i]);
      for (int it0 = 0; it0 < nitems; it0 += NST) {
        const uint32_t par = ((it0 / NST) & 1) ^ 1;
        auto stage_body = [&](auto ST) {     // stage == ST: compile-time offsets into the G ring
          constexpr int st = decltype(ST)::value;
          const int it = it0 + st;
          if (it < nitems) {
            float (&cur)[NR] = vq[st];
            if (it + NST - 1 < nitems) load_item(it + NST - 1, vq[(st + NST - 1) % NST]);
            mbar_wait(smem_u32(&bars->empty[st]), par);
#pragma unroll
            for (int rr = 0; rr < NR; ++rr)
              asm volatile("st.shared.f32 [%0+%2], %1;" ::"r"(saddr[rr]), "f"(cur[rr]), "n"(st * GST) : "memory");
            if constexpr (X3) {
#pragma unroll
              for (int rr = 0; rr < NR; ++rr)
                asm volatile("st.shared.f32 [%0+%2], %1;" ::"r"(saddr[rr]), "f"(x3_lo(cur[rr])), "n"(st * GST + GHALF)
                             : "memory");
              // X_lo from the raw X chunks TMA has landed for this item: 16 KB, two float4 per builder thread
              mbar_wait(smem_u32(&bars->x_full[st]), par ^ 1);
              uint8_t* xs = sX + (size_t)st * XSTAGE;
              const int o0 = (int)(threadIdx.x - 64) * 16;
              const float4 x0v = *reinterpret_cast<const float4*>(xs + o0);
              const float4 x1v = *reinterpret_cast<const float4*>(xs + o0 + NBUILD * 32 * 16);
              *reinterpret_cast<float4*>(xs + 4 * XCHUNK + o0) = x3_lo4(x0v);
              *reinterpret_cast<float4*>(xs + 4 * XCHUNK + o0 + NBUILD * 32 * 16) = x3_lo4(x1v);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the MMA
            __threadfence_block();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(g_full0 + (uint32_t)st * 8u);   // on the leader's barrier
          }
        };
        stage_body(std::integral_constant<int, 0>{});
        stage_body(std::integral_constant<int, 1>{});
        if constexpr (NST >= 3) stage_body(std::integral_constant<int, 2>{});
        if constexpr (NST == 4) stage_body(std::integral_constant<int, 3>{});
      }
    }

    // ---- epilogue: D^T[channel lane, output pixel column] -> out[pixel, channel] ----
    mbar_wait(smem_u32(&bars->acc_full), 0);
    tc_fence_after();
    const int qd = warp & 3;        // TMEM lane quadrant this warp may read
    const int e = bw >> 2;          // the NBUILD/4 warps of a quadrant take alternate output rows
    const float inv = P.inv;
    // One TMEM load is [32 channels (lanes)] x [32 output pixels of one row]: exactly one TMA store box.  The warp
    // writes it to shared memory as [pixel][channel] (lane = channel: conflict-free 128-byte rows) and one lane hands it
    // to the TMA unit, which also clips the columns and keeps the LSU out of the store path.  Staging: two 4 KB boxes
    // per warp in the drained X ring.
    const uint32_t stg = smem_u32(sX) + (uint32_t)bw * 8192u;
    const CUtensorMap* omap = &P.o[mode][cls];
    int nb = 0;
    for (int h = 0; h < (X3 ? 1 : P.nhalf); ++h) {
      const int c0 = (X3 ? ch : h) * 128 + qd * 32;
      if (c0 >= P.C) break;         // warp-uniform
      for (int qq = e; qq < TILE_R; qq += NBUILD / 4) {
        const int row = r0 + qq;
        if (row >= Hout) continue;  // warp-uniform
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(h * TILE_N + qq * 32), v);
        if constexpr (X3) {   // + the accumulator of the small terms, in two pieces (fewer live registers)
          float w[16];
          tmem_ld16(tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(TILE_N + qq * 32), w);
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] += w[j];
          tmem_ld16(tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(TILE_N + qq * 32 + 16), w);
#pragma unroll
          for (int j = 0; j < 16; ++j) v[16 + j] += w[j];
        }
        const uint32_t buf = stg + (uint32_t)(nb & 1) * 4096u;
        if (nb >= 2) {
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < 32; ++j)
          asm volatile("st.shared.f32 [%0], %1;" ::"r"(buf + (uint32_t)(lane * 4 + j * 128)), "f"(v[j] * inv)
                       : "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
          tma_store_4d(omap, buf, c0, x0, row, n);
          tma_store_commit();
        }
        ++nb;
      }
    }
    if (lane == 0) tma_store_wait_read<0>();   // shared memory stays valid until it was read
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();          // neither CTA may go while the other can still touch its shared memory / barriers
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, TMEM_COLS);
  }
}


constexpr int BWD_NST = 3;
constexpr int BWD_NST_X3 = 2;

}  // namespace

bool corr_tc_bwd_supported(const CorrGeom& g, int precision) {
  if (precision != FN2_PREC_TF32 && precision != FN2_PREC_TF32X3) return false;
  if (g.k != 1 || g.s1 != 1 || g.pad != g.md) return false;
  if (g.s2 != 1 && g.s2 != 2) return false;
  if (g.C % KCH != 0 || g.C / KCH > MAX_CH) return false;
  if (g.bTop % g.s2 != 0) return false;
  if (g.r != 10) return false;     // window radius instantiated below
  return true;
}

int corr_tc_backward_launch(const CorrGeom& g, const float* gout, const float* a, const float* b, float* ga, float* gb,
                            int precision, cudaStream_t st) {
  if (!corr_tc_bwd_supported(g, precision)) return FN2_E_UNSUPPORTED_PRECISION;
  const bool x3 = precision == FN2_PREC_TF32X3;
  BwdParams P;
  P.grad = gout;
  const bool fused = g.gMask != nullptr || g.gPitch != g.D;
  P.mask = g.gMask ? g.gMask : gout;   // pitch without a mask: slope 1 makes the sign test a no-op
  P.gpitch = g.gPitch;
  P.slope = g.gMask ? g.slope : 1.0f;
  P.out[0] = ga; P.out[1] = gb;
  P.N = g.N; P.H = g.H; P.W = g.W; P.C = g.C; P.s2 = g.s2; P.ncls = g.s2 * g.s2;
  P.bH = g.bH; P.boff = g.bTop / g.s2;
  for (int c = 0; c < 2; ++c) {
    P.Hc[c] = c < g.s2 ? (g.H - c + g.s2 - 1) / g.s2 : 0;
    P.Wc[c] = c < g.s2 ? (g.W - c + g.s2 - 1) / g.s2 : 0;
    P.Hcb[c] = c < g.s2 ? (g.bH - c + g.s2 - 1) / g.s2 : 0;
  }
  P.TY = (P.Hcb[0] + TILE_R - 1) / TILE_R;   // db tiles cover the (possibly taller) b-space
  P.TX = (P.Wc[0] + TILE_C - 1) / TILE_C;
  P.nchunk = g.C / KCH;
  P.nhalf = (g.C + 127) / 128;
  P.inv = 1.0f / (float)g.C;
  for (int mode = 0; mode < 2; ++mode) {
    const float* src = mode ? a : b;
    for (int cls = 0; cls < 4; ++cls) {
      const int cy = cls / g.s2, cx = cls % g.s2;
      const int srcH = mode ? g.H : g.bH;
      const int srcHc = mode ? P.Hc[cy] : P.Hcb[cy];
      if (cls >= P.ncls || srcHc < 1 || P.Wc[cx] < 1) {  // unused / empty class: any valid map, never dereferenced
        P.x[mode][cls] = P.x[mode][0];
        P.o[mode][cls] = P.o[mode][0];
        continue;
      }
      if (!make_class_map(&P.x[mode][cls], src, g.N, srcH, g.W, g.C, g.s2, cy, cx, srcHc, P.Wc[cx], TILE_C, 1,
                          CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B))
        return FN2_E_CUDA;
      float* dst = mode ? gb : ga;
      const int dstH = mode ? g.bH : g.H;
      const int dstHc = mode ? P.Hcb[cy] : P.Hc[cy];
      if (dstHc < 1) { P.o[mode][cls] = P.o[mode][0]; continue; }
      if (!make_class_map(&P.o[mode][cls], dst, g.N, dstH, g.W, g.C, g.s2, cy, cx, dstHc, P.Wc[cx], TILE_C, 1,
                          CU_TENSOR_MAP_SWIZZLE_NONE))
        return FN2_E_CUDA;
    }
  }
  using Kern = void (*)(const BwdParams);
  const Kern table[2][2][2] = {   // [x3][s2 == 2][fused]
      {{corr_tc_bwd_kernel<BWD_NST, 1, 10, false, false>, corr_tc_bwd_kernel<BWD_NST, 1, 10, false, true>},
       {corr_tc_bwd_kernel<BWD_NST, 2, 10, false, false>, corr_tc_bwd_kernel<BWD_NST, 2, 10, false, true>}},
      {{corr_tc_bwd_kernel<BWD_NST_X3, 1, 10, true, false>, corr_tc_bwd_kernel<BWD_NST_X3, 1, 10, true, true>},
       {corr_tc_bwd_kernel<BWD_NST_X3, 2, 10, true, false>, corr_tc_bwd_kernel<BWD_NST_X3, 2, 10, true, true>}}};
  const Kern kern = table[x3 ? 1 : 0][g.s2 == 2 ? 1 : 0][fused ? 1 : 0];
  const size_t smem = x3 ? bwd_smem_bytes<BWD_NST_X3, true>() : bwd_smem_bytes<BWD_NST, false>();
  static PerDeviceOnce once;   // function attributes are per device
  const cudaError_t attr_err = per_device_once(once, [&table] {
    cudaError_t e = cudaSuccess;
    for (int x = 0; x < 2 && e == cudaSuccess; ++x)
      for (int q = 0; q < 2 && e == cudaSuccess; ++q)
        for (int f = 0; f < 2 && e == cudaSuccess; ++f)
          e = cudaFuncSetAttribute(table[x][q][f], cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)(x ? bwd_smem_bytes<BWD_NST_X3, true>() : bwd_smem_bytes<BWD_NST, false>()));
    return e;
  });
  if (attr_err != cudaSuccess) { note_cuda_error(attr_err); (void)cudaGetLastError(); return FN2_E_CUDA; }
  const long long grid = (long long)g.N * P.ncls * P.TY * P.TX * 2 * (x3 ? P.nhalf : 1);
  if (grid > 0x7fffffffLL) return FN2_E_BAD_SHAPE;
  if (x3 && P.nhalf == 2 && !env_switches().no_pair) {
    // C = 256: the two channel halves of a tile as one CTA pair sharing the gradient Gram matrix
    const Kern ptable[2][2] = {{corr_tc_bwd_x3_pair_kernel<1, 10, false>, corr_tc_bwd_x3_pair_kernel<1, 10, true>},
                               {corr_tc_bwd_x3_pair_kernel<2, 10, false>, corr_tc_bwd_x3_pair_kernel<2, 10, true>}};
    static PerDeviceOnce once_pair;
    const cudaError_t pe = per_device_once(once_pair, [&ptable] {
      cudaError_t e = cudaSuccess;
      for (int q = 0; q < 2 && e == cudaSuccess; ++q)
        for (int f = 0; f < 2 && e == cudaSuccess; ++f)
          e = cudaFuncSetAttribute(ptable[q][f], cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)bwd_pair_smem_bytes());
      return e;
    });
    if (pe != cudaSuccess) { note_cuda_error(pe); (void)cudaGetLastError(); return FN2_E_CUDA; }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = bwd_pair_smem_bytes();
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, ptable[g.s2 == 2 ? 1 : 0][fused ? 1 : 0], P);
    if (e != cudaSuccess) { note_cuda_error(e); (void)cudaGetLastError(); return FN2_E_CUDA; }
    return after_launch();
  }
  kern<<<(unsigned)grid, NTHREADS, smem, st>>>(P);
  return after_launch();
}

}  // namespace fn2
