// corr_tc_x3.cu -- FlowNetC correlation FORWARD at fp32-grade accuracy on the tensor cores (split-TF32), sm_100a.
//
// The reference op is fp32 end to end (correlation_op.cc:54-61, arithmetic correlation_kernel.cu.cc:72-110).  A single
// kind::tf32 MMA keeps 11 significant bits of each operand; here every fp32 operand x is split into hi = its top 19
// bits (what the tensor core reads from the raw fp32 word anyway) and lo = x - hi (exact, rounded to 11 bits, see
// fn2_tc.cuh x3_lo), and the products hi*hi, hi*lo, lo*hi and lo*lo are all accumulated in fp32:
//   <a, b> = <a_hi, b_hi> + <a_hi, b_lo> + <a_lo, b_hi> + <a_lo, b_lo>        error ~2^-21 per product
// Same banded-Gram formulation as corr_tc.cu (parity classes, blocks of b-rows x 32 b-columns -- two rows here --, TMA
// zero fill as the op's padding), with the split folded into the MMA shape instead of into extra instructions:
//   * one CTA owns 64 a-pixels (2 class-rows x 32 class-cols).  Its A operand has M = 128 rows: rows 0..63 the raw
//     pixels, rows 64..127 their lo parts, resident in shared memory for the full K = C (128 KB at C = 256).
//   * every streamed block of 64 b-pixels (2 b-rows) becomes N = 128 operand rows: 64 raw + 64 lo.  ONE tcgen05.mma
//     M128 N128 K8 therefore produces all four partial products; they land in four quadrants of the [128 x 128]
//     accumulator:  lanes 0..63 x cols 0..63 = hi*hi, x cols 64..127 = hi*lo,
//                   lanes 64..127 x cols 0..63 = lo*hi, x cols 64..127 = lo*lo.
//     A stage of the ring is one 32-channel chunk of a block, 8 KB raw + 8 KB lo: four stages deep, so the TMA fill
//     and the conversion of the next chunks run under the MMAs of the current one (with 4-row blocks only two 32 KB
//     stages fit beside the resident a tile and the tensor core waited for every fill: 0.37 ms at N = 32).
//   * lo parts are made in shared memory by four converter warps (raw chunk -> lo chunk at the same swizzled offset),
//     so HBM / L2 traffic is that of the plain kernel: nothing is pre-split in global memory, no workspace.
//   * epilogue: a warp can only read its own TMEM lane quadrant, so the warp holding a pixel row's hi lanes and the
//     one holding its lo lanes each add their two column halves, exchange one partial through the staging tile and
//     each finishes (transpose + coalesced 441-channel store) one of the block's two b-rows.
// Cost: 4x the tensor work of the tf32 kernel per output, for an output that meets the fp32 tolerance
// (1e-5 relative + 1e-6 absolute against a float64 evaluation).
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2-5 = epilogue (one
// per TMEM lane quadrant), warps 6-9 = converters.
//
// What ships by default is the CTA-PAIR version further down (corr_tc_fwd_x3_pair_kernel: cta_group::2, M256 N256,
// persistent, narrow and wide images; 0.210 ms at N = 32 against 0.343 for the single-CTA kernel described above, which
// stays as the FN2_NO_PAIR developer fallback and as the plain statement of the scheme).
#include <algorithm>
#include "fn2_tc.cuh"

namespace fn2 {
namespace {

using namespace tc;

constexpr int XR = 2;                   // a-rows per tile (class space)
constexpr int TILE_C = 32;              // a-cols per tile
constexpr int A_HALF = 64 * 128;        // 64 a-pixels x 32 channels fp32 = 8 KB (raw; the lo copy follows it)
constexpr int A_CHUNK = 2 * A_HALF;     // M = 128 operand rows of one 32-channel chunk
constexpr int BROWS = 2;                // b-rows per block
constexpr int B_HALF = BROWS * 32 * 128;   // 64 b-pixels x 32 channels = 8 KB (raw; the lo copy follows it)
constexpr int B_STAGE = 2 * B_HALF;     // N = 128 operand rows of one 32-channel chunk
#ifndef X3_NSTAGE
#define X3_NSTAGE 4
#endif
constexpr int NSTAGE = X3_NSTAGE;       // B ring depth
constexpr int NBUF = 4;                 // TMEM accumulator buffers (4 x 128 columns = 512)
constexpr int BUF_COLS = 2 * BROWS * 32;
constexpr int MAX_KC = 8;               // C <= 256
constexpr int NEPI = 4;                 // epilogue warps: (a-row, hi/lo lanes) = one per TMEM lane quadrant
constexpr int NCONV = 4;                // converter warps
constexpr int NTHREADS = 64 + 32 * NEPI + 32 * NCONV;
constexpr int STG_FLOATS = 32 * 33;     // transposed staging tile [b column][a column], pitch 33

struct X3Params {
  CUtensorMap a[4];   // per parity class, boxes of 2 rows x 32 cols x 32 channels
  CUtensorMap b[4];   // per parity class, boxes of 2 rows x 32 cols x 32 channels
  CUtensorMap bw[4];  // per parity class, boxes of 1 row x 64 cols x 32 channels (CTA-pair kernel, wide images)
  float* out;
  int N, H, W, C, D, s2, ncls;
  int Hc[2], Wc[2];   // class-space extent per parity
  int Hcb[2];         // class-space rows of b (>= Hc + boff when b carries row halos)
  int boff;           // class-space row of b that a-row 0 meets at displacement 0
  int TY, TX;
  int TYP, nwork;     // CTA-pair kernel: pair tiles (4 a-rows) per class, length of the work list
  int nkc;            // C / 32
  float sumelems;
  int opitch;         // floats between consecutive pixels of the output
  float slope;        // fused LeakyReLU (1 = none)
};

struct __align__(8) X3Barriers {
  uint64_t a_full[MAX_KC];    // TMA: chunk kc of the raw a tile has landed
  uint64_t a_conv[MAX_KC];    // converters: its lo half is written (block 0 starts on chunk 0, not on the whole tile)
  uint64_t full[NSTAGE];      // TMA: raw b chunk landed
  uint64_t conv[NSTAGE];      // converters: lo b chunk written
  uint64_t empty[NSTAGE];     // MMAs have read the stage
  uint64_t tmem_full[NBUF];
  uint64_t tmem_empty[NBUF];
  uint32_t tmem_base;
};

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// 32 lanes x 32 columns of BOTH column halves of an accumulator buffer, summed: (x raw b) + (x lo b)
__device__ __forceinline__ void tmem_ld32_sum2(uint32_t taddr, float (&v)[32]) {
  constexpr uint32_t LO = BUF_COLS / 2;
  float w[16];   // the second half in two pieces: fewer live registers
  tmem_ld32(taddr, v);
  tmem_ld16(taddr + LO, w);
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] += w[j];
  tmem_ld16(taddr + LO + 16u, w);
#pragma unroll
  for (int j = 0; j < 16; ++j) v[16 + j] += w[j];
}

// Dynamic shared memory: [A: nkc x 16 KB][B ring: NSTAGE x 16 KB][staging: NEPI x 32x33 floats][barriers]
template <int R, bool WIDE>
__global__ void __launch_bounds__(NTHREADS, 1) corr_tc_fwd_x3_kernel(const __grid_constant__ X3Params P) {
  constexpr int GW = 2 * R + 1;
  constexpr int NHX = WIDE ? 2 : 1;                 // 32-column b blocks per group of BROWS b-rows
  constexpr int XO0 = WIDE ? -16 : 0;               // column origin of the first one, relative to the a tile
  constexpr int NTT = WIDE ? (GW * 16 + 31) / 32 : GW;  // store instructions per (a-row, b-row, block)
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = sA + (size_t)P.nkc * A_CHUNK;
  float* sStage = reinterpret_cast<float*>(sB + (size_t)NSTAGE * B_STAGE);
  X3Barriers* bars = reinterpret_cast<X3Barriers*>(sStage + NEPI * STG_FLOATS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- which tile ----
  int t = blockIdx.x;
  const int tx = t % P.TX; t /= P.TX;
  const int ty = t % P.TY; t /= P.TY;
  const int cls = t % P.ncls;
  const int n = t / P.ncls;
  const int cy = cls / P.s2, cx = cls % P.s2;
  const int Hc = P.Hc[cy], Wc = P.Wc[cx];
  const int y0 = ty * XR, x0 = tx * TILE_C;
  if (y0 >= Hc || x0 >= Wc) return;  // whole CTA leaves together, before any barrier exists
  const int Hcb = P.Hcb[cy];
  const int s_first = max(0, y0 + P.boff - R);
  const int s_last = min(Hcb - 1, y0 + P.boff + XR - 1 + R);
  const int nblk = max(0, (s_last - s_first + BROWS) / BROWS) * NHX;   // blocks of BROWS b-rows x 32 b-columns
  const int nkc = P.nkc;

  if (threadIdx.x == 0) {
    for (int i = 0; i < MAX_KC; ++i) {
      mbar_init(smem_u32(&bars->a_full[i]), 1);
      mbar_init(smem_u32(&bars->a_conv[i]), NCONV);
    }
    for (int i = 0; i < NSTAGE; ++i) {
      mbar_init(smem_u32(&bars->full[i]), 1);
      mbar_init(smem_u32(&bars->conv[i]), NCONV);
      mbar_init(smem_u32(&bars->empty[i]), 1);
    }
    for (int i = 0; i < NBUF; ++i) { mbar_init(smem_u32(&bars->tmem_full[i]), 1); mbar_init(smem_u32(&bars->tmem_empty[i]), NEPI); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(smem_u32(&bars->tmem_base), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (elect_one()) {
      for (int kc = 0; kc < nkc; ++kc) {
        const uint32_t a_full = smem_u32(&bars->a_full[kc]);
        mbar_expect_tx(a_full, A_HALF);
        tma_load_4d(smem_u32(sA + (size_t)kc * A_CHUNK), &P.a[cls], a_full, kc * KCH, x0, y0, n);
      }
      int it = 0;
      for (int blk = 0; blk < nblk; ++blk) {
        const int srow = s_first + (blk / NHX) * BROWS;
        const int xcol = x0 + XO0 + (blk % NHX) * 32;
        for (int kc = 0; kc < nkc; ++kc, ++it) {
          const int stage = it % NSTAGE;
          mbar_wait(smem_u32(&bars->empty[stage]), ((it / NSTAGE) & 1) ^ 1);
          const uint32_t full = smem_u32(&bars->full[stage]);
          mbar_expect_tx(full, B_HALF);
          tma_load_4d(smem_u32(sB + (size_t)stage * B_STAGE), &P.b[cls], full, kc * KCH, xcol, srow, n);
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    if (elect_one()) {
      constexpr uint32_t idesc = idesc_tf32(128, BUF_COLS);
      int it = 0;
      for (int blk = 0; blk < nblk; ++blk) {
        const int buf = blk % NBUF;
        mbar_wait(smem_u32(&bars->tmem_empty[buf]), ((blk / NBUF) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)buf * BUF_COLS;
        for (int kc = 0; kc < nkc; ++kc, ++it) {
          const int stage = it % NSTAGE;
          const uint32_t par = (it / NSTAGE) & 1;
          if (blk == 0) {   // the a tile arrives and is split chunk by chunk
            mbar_wait(smem_u32(&bars->a_full[kc]), 0);
            mbar_wait(smem_u32(&bars->a_conv[kc]), 0);
          }
          mbar_wait(smem_u32(&bars->full[stage]), par);
          mbar_wait(smem_u32(&bars->conv[stage]), par);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + (size_t)kc * A_CHUNK);
          const uint32_t b_addr = smem_u32(sB + (size_t)stage * B_STAGE);
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)  // K = 8 tf32 = 32 bytes per instruction inside the 128-byte swizzle row
            umma_tf32(d_tmem, desc_sw128_kmajor(a_addr + k4 * 32), desc_sw128_kmajor(b_addr + k4 * 32), idesc,
                      (uint32_t)((kc | k4) != 0));
          umma_commit(smem_u32(&bars->empty[stage]));   // frees the B stage (raw and lo) when these MMAs have read it
        }
        umma_commit(smem_u32(&bars->tmem_full[buf]));   // accumulator block complete
      }
    }
  } else if (warp < 2 + NEPI) {
    // ================================ epilogue (warps 2..5) ================================
    const int q = warp & 3;              // TMEM lane quadrant this warp may read
    const int ar = q & 1;                // a-row of the tile
    const int isLo = q >> 1;             // quadrants 2, 3 hold the products of a_lo
    // The hi warp and the lo warp of an a-row share two staging tiles: tile 0 collects b-row 0 of a block (finished
    // by the hi warp), tile 1 b-row 1 (finished by the lo warp).
    float* mine = sStage + (ar * 2 + isLo) * STG_FLOATS;
    float* theirs = sStage + (ar * 2 + (isLo ^ 1)) * STG_FLOATS;
    const int y = y0 + ar;               // class-space row of this warp's 32 pixels
    const bool row_ok = y < Hc;
    const int Y = y * P.s2 + cy;         // full-resolution row
    const int ncol = min(TILE_C, Wc - x0);
    const float inv = 1.0f / P.sumelems;
    float* orow0 = P.out + (((size_t)n * P.H + (row_ok ? Y : 0)) * P.W + (size_t)(x0 * P.s2 + cx)) * P.opitch;
    const int pixstride = P.s2 * P.opitch;    // floats between neighbouring pixels of the class

    // Per-lane plan for writing one displacement row (32 pixels x GW values, 84-byte runs) coalesced; see
    // corr_tc_fwd_kernel (corr_tc.cu) for the enumeration.
    // One packed word per store: pixel column cc [0,5), displacement o [5,10), b column xb [10,15), bit 15 = the
    // b column lies inside the block (else the value is a lateral zero), bit 16 = the store is live.
    uint32_t plan[NHX][NTT];
#pragma unroll
    for (int hx = 0; hx < NHX; ++hx) {
      const int XO = XO0 + 32 * hx;
#pragma unroll
      for (int tt = 0; tt < NTT; ++tt) {
        const int e = tt * 32 + lane;
        int cc, o;
        bool live = true;
        if (!WIDE) {
          cc = e / GW; o = e - cc * GW;
        } else {
          int acc = 0; cc = 0; o = 0; live = false;
#pragma unroll 1
          for (int c2 = 0; c2 < 32; ++c2) {
            const int lo = max(0, R + XO - c2), hi = min(GW - 1, R + XO + 31 - c2);
            const int cnt = max(0, hi - lo + 1);
            if (!live && e < acc + cnt) { cc = c2; o = lo + (e - acc); live = true; }
            acc += cnt;
          }
        }
        const int xb = cc + o - R - XO;
        const bool xok = (xb >= 0) && (xb < 32);
        plan[hx][tt] = (uint32_t)cc | ((uint32_t)o << 5) | ((uint32_t)(xok ? xb : 0) << 10) | ((xok ? 1u : 0u) << 15) |
                       (((live && cc < ncol) ? 1u : 0u) << 16);
      }
    }

    // displacement rows that point outside the image: written as zeros here, never computed
    if (row_ok) {
      const int yb = y + P.boff;                          // b-row this a-row meets at p = 0
      const int dlo = min(GW, max(0, R - yb)) * GW;       // p < -yb
      const int dhi = max(0, min(GW, Hcb - yb + R)) * GW; // p > Hcb-1-yb
      for (int cc = isLo; cc < ncol; cc += 2) {
        float* po = orow0 + (size_t)cc * pixstride;
        for (int e = lane; e < dlo; e += 32) po[e] = 0.f;
        for (int e = dhi + lane; e < P.D; e += 32) po[e] = 0.f;
      }
    }

    for (int blk0 = 0; blk0 < nblk; blk0 += NHX) {
#pragma unroll
      for (int hx = 0; hx < NHX; ++hx) {
        const int blk = blk0 + hx;
        const int buf = blk % NBUF;
        mbar_wait(smem_u32(&bars->tmem_full[buf]), (blk / NBUF) & 1);
        tc_fence_after();
        const int sbase = s_first + (blk0 / NHX) * BROWS;
        const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BUF_COLS);
        {
          const int brMine = isLo, brTheirs = isLo ^ 1;
          const int sM = sbase + brMine, sT = sbase + brTheirs;
          const int pM = sM - (y + P.boff), pT = sT - (y + P.boff);
          const bool okM = row_ok && pM >= -R && pM <= R && sM < Hcb;   // the same in both warps of the pair
          const bool okT = row_ok && pT >= -R && pT <= R && sT < Hcb;
          float vm[32];
          if (okT) {   // the partner finishes this b-row: hand it this warp's partial sum
            float v[32];
            tmem_ld32_sum2(tbase + (uint32_t)(brTheirs * 32), v);
#pragma unroll
            for (int j = 0; j < 32; ++j) theirs[j * 33 + lane] = v[j];   // [b column][a column], conflict-free
          }
          if (okM) tmem_ld32_sum2(tbase + (uint32_t)(brMine * 32), vm);
          named_bar_sync(1 + ar, 64);      // both partial sums are in the tiles
          if (okM) {
#pragma unroll
            for (int j = 0; j < 32; ++j) mine[j * 33 + lane] += vm[j];   // same lane, same word the partner wrote
            __syncwarp();
            float* orow = orow0 + (pM + R) * GW;
#pragma unroll
            for (int tt = 0; tt < NTT; ++tt) {
              const uint32_t pl = plan[hx][tt];
              const int cc = pl & 31u, o = (pl >> 5) & 31u, xb = (pl >> 10) & 31u;
              float x = mine[xb * 33 + cc];
              x = ((pl >> 15) & 1u) ? x * inv : 0.f;
              x = x > 0.f ? x : x * P.slope;
              if ((pl >> 16) & 1u) orow[cc * pixstride + o] = x;
            }
          }
          named_bar_sync(1 + ar, 64);      // the tiles may be overwritten
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bars->tmem_empty[buf]));
      }
    }
  } else {
    // ================================ converters (warps 6..9): lo = x - tf32(x) ================================
    const int ct = threadIdx.x - (64 + 32 * NEPI);
    constexpr int NCT = 32 * NCONV;
    for (int kc = 0; kc < nkc; ++kc) {
      mbar_wait(smem_u32(&bars->a_full[kc]), 0);
      uint8_t* base = sA + (size_t)kc * A_CHUNK;
      float4 v[A_HALF / 16 / NCT];
#pragma unroll
      for (int i = 0; i < A_HALF / 16 / NCT; ++i) v[i] = *reinterpret_cast<const float4*>(base + (ct + i * NCT) * 16);
#pragma unroll
      for (int i = 0; i < A_HALF / 16 / NCT; ++i)
        *reinterpret_cast<float4*>(base + A_HALF + (ct + i * NCT) * 16) = x3_lo4(v[i]);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars->a_conv[kc]));
    }
    const int nit = nblk * nkc;
    for (int it = 0; it < nit; ++it) {
      const int stage = it % NSTAGE;
      mbar_wait(smem_u32(&bars->full[stage]), (it / NSTAGE) & 1);
      uint8_t* base = sB + (size_t)stage * B_STAGE;
      float4 v[B_HALF / 16 / NCT];
#pragma unroll
      for (int i = 0; i < B_HALF / 16 / NCT; ++i) v[i] = *reinterpret_cast<const float4*>(base + (ct + i * NCT) * 16);
#pragma unroll
      for (int i = 0; i < B_HALF / 16 / NCT; ++i)
        *reinterpret_cast<float4*>(base + B_HALF + (ct + i * NCT) * 16) = x3_lo4(v[i]);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars->conv[stage]));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ---- CTA-pair version (cta_group::2), narrow images (one 32-column tile per class: the FlowNetC shape) ----------------
// A single-CTA M128 N128 K8 SS-form MMA reads 8 KB of shared memory per 64 tensor clocks, the SM's whole 128 B/clk,
// and the converters' and TMA's traffic comes on top: the kernel above is bound by shared-memory bandwidth at ~40 % of
// the tensor rate.  Here two CTAs of a cluster own FOUR consecutive class-rows of a (two each: M = 256 = 2 x [64 raw +
// 64 lo]) and share every streamed block of FOUR b-rows (N = 256 = 2 x [64 raw + 64 lo]): each CTA loads and splits
// only ITS two b-rows, and one tcgen05.mma.cta_group::2 M256 N256 K8 (128 tensor clocks) reads 4 KB of A + 4 KB of B
// per CTA -- 64 B/clk, plus 32 B/clk of conversion and 16 B/clk of TMA fill.  Accumulator: 128 lanes x 256 columns per
// CTA and block, two buffers.  Column n of a block: CTA (n / 128)'s b-rows, inside them [64 raw | 64 lo] x
// [b-row 0 | b-row 1] x 32 columns.  Persistent like corr_tc_fwd_pair_kernel (corr_tc.cu): a pair strides over the
// work list, every barrier's phase carries over, the next tile's a chunks load and split under the last block.
// Barriers: full / a_full are local (each CTA's converters wait for their own TMA); conv / a_conv live in the leader
// and collect BOTH CTAs' converter warps (the issuer's only wait before the MMAs); empty / a_free / tmem_full are
// multicast commits; tmem_empty lives in the leader and collects both CTAs' epilogue warps.
constexpr int PR = 2;                   // a-rows per CTA
constexpr int PBROWS = 4;               // b-rows per block (two per CTA)
constexpr int PNBUF = 2;                // accumulator buffers of 256 columns
constexpr int PBUF_COLS = 256;
constexpr int PNEPI = 8;                // epilogue warps: two per TMEM lane quadrant, one per CTA-half of the block
constexpr int PNTHREADS = 64 + 32 * PNEPI + 32 * NCONV;

struct __align__(8) X3PairBarriers {
  uint64_t a_full[MAX_KC];    // local: raw chunk kc of this CTA's a tile has landed
  uint64_t a_conv[MAX_KC];    // leader: both CTAs' lo halves of chunk kc are written
  uint64_t a_free[MAX_KC];    // local (multicast commit): the tile's last block has read chunk kc
  uint64_t full[NSTAGE];      // local: this CTA's raw half of the b chunk landed
  uint64_t conv[NSTAGE];      // leader: both CTAs' lo halves written
  uint64_t empty[NSTAGE];     // local (multicast commit)
  uint64_t tmem_full[PNBUF];  // local (multicast commit)
  uint64_t tmem_empty[PNBUF]; // leader: both CTAs' epilogue warps
  uint32_t tmem_base;
};

struct X3Work {
  int n, cls, cy, cx, Hc, Wc, Hcb, y0p, x0, s_first, nblk;
  bool valid;
};
template <int R, bool WIDE>
__device__ __forceinline__ X3Work x3_work(const X3Params& P, int w) {
  X3Work k;
  int t = w;
  const int tx = t % P.TX; t /= P.TX;
  const int typ = t % P.TYP; t /= P.TYP;
  k.cls = t % P.ncls;
  k.n = t / P.ncls;
  k.cy = k.cls / P.s2; k.cx = k.cls % P.s2;
  k.Hc = P.Hc[k.cy]; k.Wc = P.Wc[k.cx]; k.Hcb = P.Hcb[k.cy];
  k.y0p = typ * 2 * PR;
  k.x0 = tx * TILE_C;
  k.nblk = 0; k.s_first = 0;
  k.valid = !(k.y0p >= k.Hc || k.x0 >= k.Wc);
  if (!k.valid) return k;
  const int ylast = min(k.y0p + 2 * PR - 1, k.Hc - 1);
  k.s_first = max(0, k.y0p + P.boff - R);
  const int s_last = min(k.Hcb - 1, ylast + P.boff + R);
  constexpr int RB = WIDE ? 2 : PBROWS;               // b-rows per block
  k.nblk = max(0, (s_last - k.s_first + RB) / RB);
  return k;
}

// WIDE (class width > 32): a block is TWO b-rows of 64 columns [x0 - 16, x0 + 48) -- one row per CTA -- so that the
// +-10 window of every pixel of the 32-column a tile lies inside one block row; columns outside the image are TMA zero
// fill (the op's padding), so no displacement needs a lateral test.
template <int R, bool WIDE>
__global__ void __launch_bounds__(PNTHREADS, 1) corr_tc_fwd_x3_pair_kernel(const __grid_constant__ X3Params P) {
  constexpr int GW = 2 * R + 1;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = sA + (size_t)P.nkc * A_CHUNK;
  float* sStage = reinterpret_cast<float*>(sB + (size_t)NSTAGE * B_STAGE);
  X3PairBarriers* bars = reinterpret_cast<X3PairBarriers*>(sStage + PNEPI * STG_FLOATS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair0 = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  const int nkc = P.nkc;

  if (threadIdx.x == 0) {
    for (int i = 0; i < MAX_KC; ++i) {
      mbar_init(smem_u32(&bars->a_full[i]), 1);
      mbar_init(smem_u32(&bars->a_conv[i]), 2 * NCONV);
      mbar_init(smem_u32(&bars->a_free[i]), 1);
    }
    for (int i = 0; i < NSTAGE; ++i) {
      mbar_init(smem_u32(&bars->full[i]), 1);
      mbar_init(smem_u32(&bars->conv[i]), 2 * NCONV);
      mbar_init(smem_u32(&bars->empty[i]), 1);
    }
    for (int i = 0; i < PNBUF; ++i) {
      mbar_init(smem_u32(&bars->tmem_full[i]), 1);
      mbar_init(smem_u32(&bars->tmem_empty[i]), 2 * PNEPI);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) tmem_alloc_pair(smem_u32(&bars->tmem_base), 512);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();          // the peer's barriers exist before anything is signalled on them
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ================================ TMA producer (both CTAs, each its own rows) ================================
    if (elect_one()) {
      int it = 0, used = 0;
      for (int w = pair0; w < P.nwork; w += npairs) {
        const X3Work k = x3_work<R, WIDE>(P, w);
        if (k.nblk == 0) continue;
        const int y0 = k.y0p + (int)rank * PR;
        const int ya = y0 < k.Hc ? y0 : 0;   // a phantom second tile (odd number of row tiles) loads valid rows, unused
        for (int blk = 0; blk < k.nblk; ++blk) {
          // this CTA's part of the block: two b-rows of 32 columns, or (WIDE) one b-row of 64 columns
          const int srow = WIDE ? k.s_first + blk * 2 + (int)rank : k.s_first + blk * PBROWS + BROWS * (int)rank;
          for (int kc = 0; kc < nkc; ++kc, ++it) {
            if (blk == 0) {   // the a tile goes out chunk by chunk, interleaved with block 0 which consumes it in this order
              if (used > 0) mbar_wait(smem_u32(&bars->a_free[kc]), (used - 1) & 1);   // the previous tile is done with it
              const uint32_t a_full = smem_u32(&bars->a_full[kc]);
              mbar_expect_tx(a_full, A_HALF);
              tma_load_4d(smem_u32(sA + (size_t)kc * A_CHUNK), &P.a[k.cls], a_full, kc * KCH, k.x0, ya, k.n);
            }
            const int stage = it % NSTAGE;
            mbar_wait(smem_u32(&bars->empty[stage]), ((it / NSTAGE) & 1) ^ 1);
            const uint32_t full = smem_u32(&bars->full[stage]);
            mbar_expect_tx(full, B_HALF);
            if (WIDE)
              tma_load_4d(smem_u32(sB + (size_t)stage * B_STAGE), &P.bw[k.cls], full, kc * KCH, k.x0 - 16, srow, k.n);
            else
              tma_load_4d(smem_u32(sB + (size_t)stage * B_STAGE), &P.b[k.cls], full, kc * KCH, 0, srow, k.n);
          }
        }
        ++used;
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (leader only) ================================
    if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = idesc_tf32(256, PBUF_COLS);
      int it = 0, gb = 0, used = 0;
      for (int w = pair0; w < P.nwork; w += npairs) {
        const X3Work k = x3_work<R, WIDE>(P, w);
        if (k.nblk == 0) continue;
        for (int blk = 0; blk < k.nblk; ++blk, ++gb) {
          const int buf = gb % PNBUF;
          mbar_wait(smem_u32(&bars->tmem_empty[buf]), ((gb / PNBUF) & 1) ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + (uint32_t)buf * PBUF_COLS;
          for (int kc = 0; kc < nkc; ++kc, ++it) {
            const int stage = it % NSTAGE;
            if (blk == 0) mbar_wait(smem_u32(&bars->a_conv[kc]), used & 1);
            mbar_wait(smem_u32(&bars->conv[stage]), (it / NSTAGE) & 1);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(sA + (size_t)kc * A_CHUNK);
            const uint32_t b_addr = smem_u32(sB + (size_t)stage * B_STAGE);
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
              umma_tf32_pair(d_tmem, desc_sw128_kmajor(a_addr + k4 * 32), desc_sw128_kmajor(b_addr + k4 * 32), idesc,
                             (uint32_t)((kc | k4) != 0));
            umma_commit_pair(smem_u32(&bars->empty[stage]), 3);   // frees the stage (raw and lo) in both CTAs
            if (blk == k.nblk - 1) umma_commit_pair(smem_u32(&bars->a_free[kc]), 3);   // a chunk kc may be replaced
          }
          umma_commit_pair(smem_u32(&bars->tmem_full[buf]), 3);   // accumulator block complete in both CTAs
        }
        ++used;
      }
    }
  } else if (warp < 2 + PNEPI) {
    // ================================ epilogue (warps 2..9, both CTAs) ================================
    const int ew = warp - 2;
    const int q = warp & 3;              // TMEM lane quadrant this warp may read
    const int ar = q & 1;                // a-row of this CTA's tile
    const int isLo = q >> 1;             // quadrants 2, 3 hold the products of a_lo
    const int half = ew >> 2;            // which CTA's b-rows of a block (columns half * 128 ...) this warp works on
    // The hi warp and the lo warp of (a-row, half) share two staging tiles: tile 0 collects b-row 0 of the half
    // (finished by the hi warp), tile 1 b-row 1 (finished by the lo warp).
    const int grp = ar * 2 + half;
    float* mine = sStage + (grp * 2 + isLo) * STG_FLOATS;
    float* theirs = sStage + (grp * 2 + (isLo ^ 1)) * STG_FLOATS;
    const float inv = 1.0f / P.sumelems;
    const int pixstride = P.s2 * P.opitch;
    const uint32_t tmem_empty0 = mapa_u32(smem_u32(&bars->tmem_empty[0]), 0);

    // Store plan of one displacement row (32 pixels x GW values, 84-byte runs, written coalesced): element e =
    // tt * 32 + lane is value o = e % GW of pixel column cc = e / GW and comes from b column xb = cc + o - R (a zero
    // when that column is outside the class row).  Recomputed per store: a table of GW registers costs more here.
    int gb = 0;
    for (int w = pair0; w < P.nwork; w += npairs) {
      const X3Work k = x3_work<R, WIDE>(P, w);
      if (!k.valid) continue;
      const int y = k.y0p + (int)rank * PR + ar;
      const bool row_ok = y < k.Hc;
      const int Y = y * P.s2 + k.cy;
      const int ncol = min(TILE_C, k.Wc - k.x0);
      float* orow0 = P.out + (((size_t)k.n * P.H + (row_ok ? Y : 0)) * P.W + (size_t)(k.x0 * P.s2 + k.cx)) * P.opitch;

      if (row_ok) {   // displacement rows that point outside the image: zeros, never computed
        const int yb = y + P.boff;
        const int dlo = min(GW, max(0, R - yb)) * GW;
        const int dhi = max(0, min(GW, k.Hcb - yb + R)) * GW;
        for (int cc = isLo + 2 * half; cc < ncol; cc += 4) {
          float* po = orow0 + (size_t)cc * pixstride;
          for (int e = lane; e < dlo; e += 32) po[e] = 0.f;
          for (int e = dhi + lane; e < P.D; e += 32) po[e] = 0.f;
        }
      }

      for (int blk = 0; blk < k.nblk; ++blk, ++gb) {
        const int buf = gb % PNBUF;
        mbar_wait(smem_u32(&bars->tmem_full[buf]), (gb / PNBUF) & 1);
        tc_fence_after();
        if constexpr (WIDE) {
          // this (a-row, half) pair of warps finishes b-row s of the block: 64 columns, tile 0 = columns 0..31 (the lo
          // warp's partial lands there first), tile 1 = columns 32..63 (the hi warp's)
          const int sR = k.s_first + blk * 2 + half;
          const int pR = sR - (y + P.boff);
          const bool ok = row_ok && pR >= -R && pR <= R && sR < k.Hcb;       // the same in both warps of the pair
          const uint32_t tb = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * PBUF_COLS + half * 128);
          float* t0 = sStage + (grp * 2) * STG_FLOATS;
          float* t1 = t0 + STG_FLOATS;
          float vm[32];
          if (ok) {
            float v[32];
            tmem_ld32_sum2(tb + (isLo ? 0u : 32u), v);                       // the half the PARTNER completes
            float* dst = isLo ? t0 : t1;
#pragma unroll
            for (int j = 0; j < 32; ++j) dst[j * 33 + lane] = v[j];
            tmem_ld32_sum2(tb + (isLo ? 32u : 0u), vm);                      // the half this warp completes
          }
          named_bar_sync(1 + grp, 64);
          if (ok) {
            float* own = isLo ? t1 : t0;
#pragma unroll
            for (int j = 0; j < 32; ++j) own[j * 33 + lane] += vm[j];
          }
          named_bar_sync(1 + grp, 64);       // both tiles hold the complete sums
          if (ok) {
            // each warp stores 16 of the 32 pixel columns: 16 x GW values, coalesced
            float* orow = orow0 + (pR + R) * GW;
            const int c16 = isLo * 16;
            constexpr int NE = 16 * GW;
#pragma unroll
            for (int t0i = 0; t0i < (NE + 31) / 32; t0i += 4) {
              float xv[4];
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                const int e = min((t0i + u) * 32 + lane, NE - 1);
                const int cc = c16 + e / GW, xb = cc + (e - (e / GW) * GW) - R + 16;
                xv[u] = t0[(xb >> 5) * STG_FLOATS + (xb & 31) * 33 + cc];
              }
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                const int e = (t0i + u) * 32 + lane;
                const int cc = c16 + e / GW, o = e - (e / GW) * GW;
                float x = xv[u] * inv;
                x = x > 0.f ? x : x * P.slope;
                if (e < NE && cc < ncol) orow[cc * pixstride + o] = x;
              }
            }
          }
          named_bar_sync(1 + grp, 64);       // the tiles may be overwritten
        } else {
        const int sbase = k.s_first + blk * PBROWS + BROWS * half;
          const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * PBUF_COLS + half * 128);
          {
            const int brMine = isLo, brTheirs = isLo ^ 1;
            const int sM = sbase + brMine, sT = sbase + brTheirs;
            const int pM = sM - (y + P.boff), pT = sT - (y + P.boff);
            const bool okM = row_ok && pM >= -R && pM <= R && sM < k.Hcb;   // the same in both warps of the pair
            const bool okT = row_ok && pT >= -R && pT <= R && sT < k.Hcb;
            float vm[32];
            if (okT) {   // the partner finishes this b-row: hand it this warp's partial sum
              float v[32];
              tmem_ld32_sum2(tbase + (uint32_t)(brTheirs * 32), v);
#pragma unroll
              for (int j = 0; j < 32; ++j) theirs[j * 33 + lane] = v[j];   // [b column][a column], conflict-free
            }
            if (okM) tmem_ld32_sum2(tbase + (uint32_t)(brMine * 32), vm);
            named_bar_sync(1 + grp, 64);     // both partial sums are in the tiles
            if (okM) {
#pragma unroll
              for (int j = 0; j < 32; ++j) mine[j * 33 + lane] += vm[j];   // same lane, same word the partner wrote
              __syncwarp();
              float* orow = orow0 + (pM + R) * GW;
              // loads in batches of 7 ahead of their stores: a load -> scale -> store chain per value left the warp
              // waiting on shared-memory latency 21 times per row (48 % of the epilogue's stall samples)
#pragma unroll
              for (int t0 = 0; t0 < GW; t0 += 7) {
                float xv[7];
#pragma unroll
                for (int u = 0; u < 7; ++u) {
                  const int e = (t0 + u) * 32 + lane;
                  const int cc = e / GW, xb = cc + (e - cc * GW) - R;
                  xv[u] = mine[(xb & 31) * 33 + cc];
                }
#pragma unroll
                for (int u = 0; u < 7; ++u) {
                  const int e = (t0 + u) * 32 + lane;
                  const int cc = e / GW, o = e - cc * GW;
                  const int xb = cc + o - R;
                  float x = (xb >= 0 && xb < 32) ? xv[u] * inv : 0.f;
                  x = x > 0.f ? x : x * P.slope;
                  if (cc < ncol) orow[cc * pixstride + o] = x;
                }
              }
            }
            named_bar_sync(1 + grp, 64);     // the tiles may be overwritten
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(tmem_empty0 + (uint32_t)buf * 8u);
      }
    }
  } else {
    // ================================ converters (warps 10..13, both CTAs): lo = x - tf32(x) ====================
    const int ct = threadIdx.x - (64 + 32 * PNEPI);
    constexpr int NCT = 32 * NCONV;
    const uint32_t a_conv0 = mapa_u32(smem_u32(&bars->a_conv[0]), 0);
    const uint32_t conv0 = mapa_u32(smem_u32(&bars->conv[0]), 0);
    int it = 0, used = 0;
    for (int w = pair0; w < P.nwork; w += npairs) {
      const X3Work k = x3_work<R, WIDE>(P, w);
      if (k.nblk == 0) continue;
      for (int blk = 0; blk < k.nblk; ++blk) {
        for (int kc = 0; kc < nkc; ++kc, ++it) {
          if (blk == 0) {
            mbar_wait(smem_u32(&bars->a_full[kc]), used & 1);
            uint8_t* base = sA + (size_t)kc * A_CHUNK;
            float4 v[A_HALF / 16 / NCT];
#pragma unroll
            for (int i = 0; i < A_HALF / 16 / NCT; ++i) v[i] = *reinterpret_cast<const float4*>(base + (ct + i * NCT) * 16);
#pragma unroll
            for (int i = 0; i < A_HALF / 16 / NCT; ++i)
              *reinterpret_cast<float4*>(base + A_HALF + (ct + i * NCT) * 16) = x3_lo4(v[i]);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA
            __threadfence_block();   // ... and ordered before the arrival below (a release at cluster scope costs a
            __syncwarp();            // MEMBAR.GPU + ERRBAR per arrival: measured 16 % of all stall samples)
            if (lane == 0) mbar_arrive_cluster(a_conv0 + (uint32_t)kc * 8u);
          }
          const int stage = it % NSTAGE;
          mbar_wait(smem_u32(&bars->full[stage]), (it / NSTAGE) & 1);
          uint8_t* base = sB + (size_t)stage * B_STAGE;
          float4 v[B_HALF / 16 / NCT];
#pragma unroll
          for (int i = 0; i < B_HALF / 16 / NCT; ++i) v[i] = *reinterpret_cast<const float4*>(base + (ct + i * NCT) * 16);
#pragma unroll
          for (int i = 0; i < B_HALF / 16 / NCT; ++i)
            *reinterpret_cast<float4*>(base + B_HALF + (ct + i * NCT) * 16) = x3_lo4(v[i]);
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __threadfence_block();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(conv0 + (uint32_t)stage * 8u);
        }
      }
      ++used;
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();          // neither CTA may go while the other can still touch its shared memory / barriers
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

size_t x3_pair_smem_bytes(int nkc) {
  return 1024 + (size_t)nkc * A_CHUNK + (size_t)NSTAGE * B_STAGE + (size_t)PNEPI * STG_FLOATS * sizeof(float) +
         sizeof(X3PairBarriers);
}

size_t x3_smem_bytes(int nkc) {
  return 1024 + (size_t)nkc * A_CHUNK + (size_t)NSTAGE * B_STAGE + (size_t)NEPI * STG_FLOATS * sizeof(float) +
         sizeof(X3Barriers);
}

}  // namespace

int corr_tc_x3_forward(const CorrGeom& g, const float* a, const float* b, float* out, cudaStream_t st) {
  X3Params P;
  P.out = out;
  P.N = g.N; P.H = g.H; P.W = g.W; P.C = g.C; P.D = g.D; P.s2 = g.s2; P.ncls = g.s2 * g.s2;
  P.boff = g.bTop / g.s2;
  for (int c = 0; c < 2; ++c) {
    P.Hc[c] = c < g.s2 ? (g.H - c + g.s2 - 1) / g.s2 : 0;
    P.Wc[c] = c < g.s2 ? (g.W - c + g.s2 - 1) / g.s2 : 0;
    P.Hcb[c] = c < g.s2 ? (g.bH - c + g.s2 - 1) / g.s2 : 0;
  }
  P.TY = (P.Hc[0] + XR - 1) / XR;
  P.TX = (P.Wc[0] + TILE_C - 1) / TILE_C;
  P.nkc = g.C / KCH;
  P.sumelems = (float)g.C;
  P.opitch = g.oPitch;
  P.slope = g.slope;
  for (int cls = 0; cls < P.ncls; ++cls) {
    const int cy = cls / g.s2, cx = cls % g.s2;
    if (P.Hc[cy] < 1 || P.Wc[cx] < 1) {  // empty class (H or W == 1 with s2 == 2): map class 0 instead, never used
      P.a[cls] = P.a[0]; P.b[cls] = P.b[0]; P.bw[cls] = P.bw[0];
      continue;
    }
    if (!make_class_map(&P.a[cls], a, g.N, g.H, g.W, g.C, g.s2, cy, cx, P.Hc[cy], P.Wc[cx], TILE_C, XR) ||
        !make_class_map(&P.b[cls], b, g.N, g.bH, g.W, g.C, g.s2, cy, cx, P.Hcb[cy], P.Wc[cx], TILE_C, BROWS) ||
        !make_class_map(&P.bw[cls], b, g.N, g.bH, g.W, g.C, g.s2, cy, cx, P.Hcb[cy], P.Wc[cx], 2 * TILE_C, 1))
      return FN2_E_CUDA;
  }
  for (int cls = P.ncls; cls < 4; ++cls) { P.a[cls] = P.a[0]; P.b[cls] = P.b[0]; P.bw[cls] = P.bw[0]; }
  static PerDeviceOnce once;
  const cudaError_t attr_err = per_device_once(once, [] {
    cudaError_t e = cudaFuncSetAttribute(corr_tc_fwd_x3_kernel<10, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)x3_smem_bytes(MAX_KC));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(corr_tc_fwd_x3_kernel<10, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)x3_smem_bytes(MAX_KC));
    return e;
  });
  if (attr_err != cudaSuccess) { note_cuda_error(attr_err); (void)cudaGetLastError(); return FN2_E_CUDA; }
  P.TYP = (P.Hc[0] + 2 * PR - 1) / (2 * PR);
  P.nwork = 0;
  static PerDeviceOnce once_pair;
  const EnvSwitches& sw = env_switches();
  if (!sw.no_pair) {
    // persistent CTA pairs over the list of pair tiles (4 a-rows x 32 columns)
    const cudaError_t pe = per_device_once(once_pair, [] {
      cudaError_t e = cudaFuncSetAttribute(corr_tc_fwd_x3_pair_kernel<10, false>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)x3_pair_smem_bytes(MAX_KC));
      if (e == cudaSuccess)
        e = cudaFuncSetAttribute(corr_tc_fwd_x3_pair_kernel<10, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)x3_pair_smem_bytes(MAX_KC));
      return e;
    });
    if (pe != cudaSuccess) { note_cuda_error(pe); (void)cudaGetLastError(); return FN2_E_CUDA; }
    const long long nwork = (long long)g.N * P.ncls * P.TYP * P.TX;
    if (nwork > 0x3fffffffLL) return FN2_E_BAD_SHAPE;
    P.nwork = (int)nwork;
    const long long slots = std::max(1, device_sm_count() / 2);
    const long long pgrid = 2 * (sw.no_persist ? nwork : std::min<long long>(nwork, slots));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)pgrid);
    cfg.blockDim = dim3(PNTHREADS);
    cfg.dynamicSmemBytes = x3_pair_smem_bytes(P.nkc);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t e = P.TX > 1 ? cudaLaunchKernelEx(&cfg, corr_tc_fwd_x3_pair_kernel<10, true>, P)
                                   : cudaLaunchKernelEx(&cfg, corr_tc_fwd_x3_pair_kernel<10, false>, P);
    if (e != cudaSuccess) { note_cuda_error(e); (void)cudaGetLastError(); return FN2_E_CUDA; }
    return after_launch();
  }
  const long long grid = (long long)g.N * P.ncls * P.TY * P.TX;
  if (grid > 0x7fffffffLL) return FN2_E_BAD_SHAPE;
  const size_t smem = x3_smem_bytes(P.nkc);
  if (P.TX > 1)
    corr_tc_fwd_x3_kernel<10, true><<<(unsigned)grid, NTHREADS, smem, st>>>(P);
  else
    corr_tc_fwd_x3_kernel<10, false><<<(unsigned)grid, NTHREADS, smem, st>>>(P);
  return after_launch();
}

}  // namespace fn2
