// corr_tc.cu -- FlowNetC correlation on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), sm_100a.
//
// Formulation (kernel_size 1, stride_1 1, pad == max_displacement; the FlowNetC call, flownet_c.py:40):
//   out[n,y,x,(p,o)] = 1/C * sum_c a[n,y,x,c] * b[n,y+s2*p,x+s2*o,c]
// With stride_2 = 2 the displacements are even, so pixels only ever meet pixels of the same (y%2, x%2) parity class.
// Inside one class the op is a stride-1 window correlation on the half-resolution grid: a banded Gram matrix
//   S[pixel i of a, pixel j of b] = <a_i, b_j>,  needed where |yj-yi| <= r and |xj-xi| <= r.
// One CTA owns an M = 128 tile of a-pixels (4 class-rows x 32 class-cols), keeps it resident in shared memory for the
// full K = C (fp32, read by the tensor core as tf32: no conversion pass), and streams the b-rows it needs through a
// TMA ring in blocks of 128 b-pixels x 32 channels.  Each block is one [128 x 128] accumulator in TMEM (4 buffers of
// 128 columns), produced by C/8 tcgen05.mma (M128 N128 K8, kind::tf32) and drained by four epilogue warps that pick
// the +-r window out of each TMEM row, restage it through shared memory and write the 441-channel output coalesced.
// Zero padding is TMA out-of-bounds fill; b-rows outside the image are never loaded, their outputs are written as 0.
//
// Wide images (class width > 32, e.g. the 4K strips of cfg5): column tiles of 32 a-pixels need b-columns
// x0-10 .. x0+41, so every group of 4 b-rows is visited as TWO 32-column blocks with origins x0-16 and x0+16 (negative
// and past-the-edge coordinates are TMA zero fill = the op's padding); each (pixel, displacement) pair lives in
// exactly one of the two, and each half has its own per-lane store plan.
// Row halos (strip partitioning, fn2_correlation_forward_halo): b may be a taller tensor than a; a-row y meets b-rows
// y + boff + p in class space, rows outside b are the op's zero padding.
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2-9 = epilogue
// (two warps per TMEM lane quadrant, taking alternate b-rows).
#include <algorithm>
#include <cstdlib>
#include "fn2_tc.cuh"

namespace fn2 {
namespace {

using namespace tc;

constexpr int TILE_R = 4;          // a-rows per tile (class space)
constexpr int TILE_C = 32;         // a-cols per tile
constexpr int BLK_N = 128;         // b-pixels per accumulator block
constexpr int CHUNK_BYTES = 128 * 128;  // 128 rows x 128 B
constexpr int NSTAGE = 4;          // B ring depth
constexpr int NBUF = 4;            // TMEM accumulator buffers (4 x 128 columns = 512)
constexpr int MAX_KC = 8;          // C <= 256
constexpr int NEPI = 8;            // epilogue warps: two per TMEM lane quadrant
constexpr int NTHREADS = 64 + 32 * NEPI;
constexpr int STG_FLOATS = 32 * 33;  // per-warp transposed staging tile [b column][a column], pitch 33

struct TcParams {
  CUtensorMap a[4];
  CUtensorMap b[4];
  CUtensorMap b2[4];  // the same tensor with boxes of 2 rows (CTA-pair kernel: half a block per CTA)
  float* out;
  int N, H, W, C, D, s2, ncls;
  int Hc[2], Wc[2];   // class-space extent per parity
  int Hcb[2];         // class-space rows of b (>= Hc + boff when b carries row halos)
  int boff;           // class-space row of b that a-row 0 meets at displacement 0 (halo_top / s2)
  int TY, TX;         // tiles per class (upper bound over classes)
  int TYP;            // pairs of row tiles per class (CTA-pair kernel)
  // Tail split (CTA-pair kernel): pair tiles >= n_full -- the ones that would run alone in a last, mostly empty wave --
  // are cut into split_units pieces of split_g b blocks each; every piece loads the a tile again.
  int n_full, split_units, split_g;
  int nwork;          // pair tiles + pieces (CTA-pair kernel: the work list a persistent pair strides over)
  int nkc;            // C / 32
  float sumelems;
  int opitch;         // floats between consecutive pixels of the output (D, or the width of a concat buffer)
  float slope;        // fused LeakyReLU: values <= 0 are multiplied by this (1 = plain correlation)
};

__device__ __forceinline__ float leaky(float x, float slope) { return x > 0.f ? x : x * slope; }

struct __align__(8) Barriers {
  uint64_t a_full;
  uint64_t full[NSTAGE];
  uint64_t empty[NSTAGE];
  uint64_t tmem_full[NBUF];
  uint64_t tmem_empty[NBUF];
  uint32_t tmem_base;
};


// Dynamic shared memory: [A: nkc x 16 KB][B ring: NSTAGE x 16 KB][staging: NEPI x 32x33 floats][barriers]
template <int R, bool WIDE>
__global__ void __launch_bounds__(NTHREADS, 1) corr_tc_fwd_kernel(const __grid_constant__ TcParams P) {
  constexpr int GW = 2 * R + 1;
  constexpr int NHX = WIDE ? 2 : 1;                 // 32-column b blocks per group of 4 b-rows
  constexpr int XO0 = WIDE ? -16 : 0;               // column origin of the first one, relative to the a tile
  constexpr int NTT = WIDE ? (GW * 16 + 31) / 32 : GW;  // store instructions per (a-row, b-row, block)
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // keep shared-space provenance (plain LDS/STS, not generic LD/ST): index the array instead of casting through ints
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = sA + (size_t)P.nkc * CHUNK_BYTES;
  float* sStage = reinterpret_cast<float*>(sB + (size_t)NSTAGE * CHUNK_BYTES);
  Barriers* bars = reinterpret_cast<Barriers*>(sStage + NEPI * STG_FLOATS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- which tile ----
  int t = blockIdx.x;
  const int tx = t % P.TX; t /= P.TX;
  const int ty = t % P.TY; t /= P.TY;
  const int cls = t % P.ncls;
  const int n = t / P.ncls;
  const int cy = cls / P.s2, cx = cls % P.s2;
  const int Hc = P.Hc[cy], Wc = P.Wc[cx];
  const int y0 = ty * TILE_R, x0 = tx * TILE_C;
  if (y0 >= Hc || x0 >= Wc) return;  // whole CTA leaves together, before any barrier exists
  const int Hcb = P.Hcb[cy];
  const int s_first = max(0, y0 + P.boff - R);
  const int s_last = min(Hcb - 1, y0 + P.boff + TILE_R - 1 + R);
  const int nblk = max(0, (s_last - s_first + 1 + 3) / 4) * NHX;   // blocks of 4 b-rows x 32 b-columns
  const int nkc = P.nkc;

  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bars->a_full), 1);
    for (int i = 0; i < NSTAGE; ++i) { mbar_init(smem_u32(&bars->full[i]), 1); mbar_init(smem_u32(&bars->empty[i]), 1); }
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
      const uint32_t a_full = smem_u32(&bars->a_full);
      mbar_expect_tx(a_full, (uint32_t)nkc * CHUNK_BYTES);
      for (int kc = 0; kc < nkc; ++kc)
        tma_load_4d(smem_u32(sA + (size_t)kc * CHUNK_BYTES), &P.a[cls], a_full, kc * KCH, x0, y0, n);
      int it = 0;
      for (int blk = 0; blk < nblk; ++blk) {
        const int srow = s_first + (blk / NHX) * 4;
        const int xcol = x0 + XO0 + (blk % NHX) * 32;
        for (int kc = 0; kc < nkc; ++kc, ++it) {
          const int stage = it % NSTAGE;
          mbar_wait(smem_u32(&bars->empty[stage]), ((it / NSTAGE) & 1) ^ 1);
          const uint32_t full = smem_u32(&bars->full[stage]);
          mbar_expect_tx(full, CHUNK_BYTES);
          tma_load_4d(smem_u32(sB + (size_t)stage * CHUNK_BYTES), &P.b[cls], full, kc * KCH, xcol, srow, n);
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    if (elect_one()) {
      constexpr uint32_t idesc = idesc_tf32(128, BLK_N);
      mbar_wait(smem_u32(&bars->a_full), 0);
      tc_fence_after();
      int it = 0;
      for (int blk = 0; blk < nblk; ++blk) {
        const int buf = blk % NBUF;
        mbar_wait(smem_u32(&bars->tmem_empty[buf]), ((blk / NBUF) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)buf * BLK_N;
        for (int kc = 0; kc < nkc; ++kc, ++it) {
          const int stage = it % NSTAGE;
          mbar_wait(smem_u32(&bars->full[stage]), (it / NSTAGE) & 1);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + (size_t)kc * CHUNK_BYTES);
          const uint32_t b_addr = smem_u32(sB + (size_t)stage * CHUNK_BYTES);
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)  // K = 8 tf32 = 32 bytes per instruction inside the 128-byte swizzle row
            umma_tf32(d_tmem, desc_sw128_kmajor(a_addr + k4 * 32), desc_sw128_kmajor(b_addr + k4 * 32), idesc,
                      (uint32_t)((kc | k4) != 0));
          umma_commit(smem_u32(&bars->empty[stage]));   // frees the B stage when these MMAs have read it
        }
        umma_commit(smem_u32(&bars->tmem_full[buf]));   // accumulator block complete
      }
    }
  } else {
    // ================================ epilogue (warps 2..9) ================================
    const int ew = warp - 2;
    const int q = warp & 3;              // TMEM lane quadrant this warp may read = a-row of the tile
    const int half = ew >> 2;            // the two warps of a quadrant take alternate b-rows
    const int y = y0 + q;                // class-space row of this warp's 32 pixels
    float* stg = sStage + ew * STG_FLOATS;
    const bool row_ok = y < Hc;
    const int Y = y * P.s2 + cy;         // full-resolution row
    const int ncol = min(TILE_C, Wc - x0);
    const float inv = 1.0f / P.sumelems;
    // first output element of this warp's pixel row (pixel column x0 of the class)
    float* orow0 = P.out + (((size_t)n * P.H + (row_ok ? Y : 0)) * P.W + (size_t)(x0 * P.s2 + cx)) * P.opitch;
    const int pixstride = P.s2 * P.opitch;    // floats between neighbouring pixels of the class

    // Per-lane plan for writing one displacement row (32 pixels x GW values, 84-byte runs) coalesced.
    // Block hx covers b-columns [XO, XO + 32) relative to the tile (XO = XO0 + 32 * hx); pixel cc keeps from it the
    // displacements o with 0 <= cc + o - R - XO < 32.  The kept (cc, o) pairs are enumerated pixel-major and dealt to
    // the lanes 32 at a time: element e = tt*32 + lane.  !WIDE: every pair is enumerated (the ones whose b-column
    // falls outside the single block are outside the image: written as zeros).
    int rd[NHX][NTT], gl[NHX][NTT];
    uint32_t xmask = 0, smask[NHX];
#pragma unroll
    for (int hx = 0; hx < NHX; ++hx) {
      smask[hx] = 0;
      const int XO = XO0 + 32 * hx;
#pragma unroll
      for (int tt = 0; tt < NTT; ++tt) {
        const int e = tt * 32 + lane;
        int cc, o;
        bool live = true;
        if (!WIDE) {
          cc = e / GW; o = e - cc * GW;
        } else {
          // pixel cc keeps o in [lo, hi]; find the pixel whose run contains element e
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
        rd[hx][tt] = xok ? xb * 33 + cc : 0;
        gl[hx][tt] = cc * pixstride + o;
        if (hx == 0) xmask |= (xok ? 1u : 0u) << tt;
        smask[hx] |= ((live && cc < ncol) ? 1u : 0u) << tt;
      }
    }

    // displacement rows that point outside the image: written as zeros here, never computed
    if (row_ok) {
      const int yb = y + P.boff;                          // b-row this a-row meets at p = 0
      const int dlo = min(GW, max(0, R - yb)) * GW;       // p < -yb
      const int dhi = max(0, min(GW, Hcb - yb + R)) * GW; // p > Hcb-1-yb
      for (int cc = half; cc < ncol; cc += 2) {
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
#pragma unroll
        for (int h2 = 0; h2 < 2; ++h2) {
          const int br = half + 2 * h2;
          const int s = s_first + (blk0 / NHX) * 4 + br;
          const int p = s - (y + P.boff);
          if (row_ok && p >= -R && p <= R && s < Hcb) {   // warp-uniform
            float v[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BLK_N + br * 32), v);
#pragma unroll
            for (int j = 0; j < 32; ++j) stg[j * 33 + lane] = v[j];   // [b column][a column], conflict-free
            __syncwarp();
            float* orow = orow0 + (p + R) * GW;
#pragma unroll
            for (int tt = 0; tt < NTT; ++tt) {
              float x = stg[rd[hx][tt]];
              x = (WIDE || ((xmask >> tt) & 1u)) ? x * inv : 0.f;
              x = x > 0.f ? x : x * P.slope;
              if ((smask[hx] >> tt) & 1u) orow[gl[hx][tt]] = x;
            }
            __syncwarp();
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bars->tmem_empty[buf]));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ---- CTA-pair forward (cta_group::2), narrow images (one column tile per class) ---------------------------------
// Two CTAs of a cluster own 8 consecutive class-rows of a (4 each, resident as before) and share every streamed b
// block: each CTA loads only HALF of a block (2 b-rows x 32 columns x 32 channels = 8 KB per chunk) and one
// tcgen05.mma.cta_group::2 (M 256, N 128, K 8) multiplies both a tiles by the whole block, reading the two halves from
// the two CTAs' shared memory.  Per a tile the L2 -> shared-memory fill drops from <= 896 KB to <= 576 KB (the pair
// needs b-rows y0-10 .. y0+17: 7 blocks of 64 KB per CTA instead of 6 of 128 KB), which is what bounds this kernel.
// Barrier protocol: the leader (rank 0) owns full[] (both CTAs' TMA bytes land on it) and tmem_empty[] (both CTAs'
// epilogue warps arrive on it); empty[] and tmem_full[] live in each CTA and are signalled by multicast commits.
#ifdef FN2_TIMELINE
__device__ long long g_tl[1024 * 64];
#define TL(slot) do { if (blockIdx.x < 1024) g_tl[blockIdx.x * 64 + (slot)] = clock64(); } while (0)
#define TLG(slot) do { if (blockIdx.x < 1024) { unsigned long long gt_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_)); g_tl[blockIdx.x * 64 + (slot)] = (long long)gt_; } } while (0)
#else
#define TL(slot) do {} while (0)
#define TLG(slot) do {} while (0)
#endif
constexpr int PSTAGE_BYTES = 64 * 128;   // half a block per CTA and chunk
constexpr int PNSTAGE = 8;

struct __align__(8) PairBarriers {
  uint64_t a_chunk[MAX_KC];   // leader: 32 channels of BOTH CTAs' a tiles have landed (block 0 starts on chunk 0)
  uint64_t a_free[MAX_KC];    // each CTA: the last block of the tile has read chunk kc of the a tile
  uint64_t full[PNSTAGE];
  uint64_t empty[PNSTAGE];
  uint64_t tmem_full[NBUF];
  uint64_t tmem_empty[NBUF];
  uint32_t tmem_base;
};

// One unit of work of the pair kernel: a pair tile (8 class-rows of a), or a piece of one (tail split).
struct PairWork {
  int n, cls, cy, cx, Hc, Wc, Hcb, y0p, s_first, nblk;
  bool first_part, valid;
};

template <int R>
__device__ __forceinline__ PairWork pair_work(const TcParams& P, int w) {
  PairWork k;
  int t = w, part = -1;
  if (t >= P.n_full) {
    const int u = t - P.n_full;
    t = P.n_full + u / P.split_units;
    part = u % P.split_units;
  }
  const int typ = t % P.TYP; t /= P.TYP;
  k.cls = t % P.ncls;
  k.n = t / P.ncls;
  k.cy = k.cls / P.s2; k.cx = k.cls % P.s2;
  k.Hc = P.Hc[k.cy]; k.Wc = P.Wc[k.cx]; k.Hcb = P.Hcb[k.cy];
  k.y0p = typ * 2 * TILE_R;
  k.first_part = part <= 0;                            // the piece that also writes the out-of-image zeros
  k.nblk = 0; k.s_first = 0;
  k.valid = !(k.y0p >= k.Hc || k.Wc < 1);
  if (!k.valid) return k;                              // this class has fewer row tiles: nothing to do
  const int ylast = min(k.y0p + 2 * TILE_R - 1, k.Hc - 1);
  k.s_first = max(0, k.y0p + P.boff - R);
  const int s_last = min(k.Hcb - 1, ylast + P.boff + R);
  k.nblk = max(0, (s_last - k.s_first + 1 + 3) / 4);   // blocks of 4 b-rows x 32 b-columns, the same in both CTAs
  if (part >= 0) {
    const int lo = part * P.split_g;
    k.s_first += lo * 4;
    k.nblk = max(0, min(P.split_g, k.nblk - lo));
    if (k.nblk == 0 && part > 0) k.valid = false;       // this tile has fewer blocks than pieces
  }
  return k;
}

// Persistent: a pair walks the work list w = pair index, + number of pairs, ...  The three roles keep running
// counters (ring slot, accumulator buffer, tile) so every barrier's phase carries over from tile to tile; the only
// tile-to-tile hazard is the resident a tile, whose chunk kc may be overwritten once the LAST block's MMAs of that
// chunk have completed (a_free[kc], committed by the issuer to both CTAs).  The next tile's a chunks and first b
// blocks therefore load under the current tile's last block and epilogue, and the prologue is paid once per CTA.
template <int R>
__global__ void __launch_bounds__(NTHREADS, 1) corr_tc_fwd_pair_kernel(const __grid_constant__ TcParams P) {
  constexpr int GW = 2 * R + 1;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = sA + (size_t)P.nkc * CHUNK_BYTES;
  float* sStage = reinterpret_cast<float*>(sB + (size_t)PNSTAGE * PSTAGE_BYTES);
  PairBarriers* bars = reinterpret_cast<PairBarriers*>(sStage + NEPI * STG_FLOATS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair0 = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  const int nkc = P.nkc;
  constexpr int x0 = 0;
  if (threadIdx.x == 0) { TL(0); TLG(1); }

  if (threadIdx.x == 0) {
    for (int i = 0; i < MAX_KC; ++i) { mbar_init(smem_u32(&bars->a_chunk[i]), 1); mbar_init(smem_u32(&bars->a_free[i]), 1); }
    for (int i = 0; i < PNSTAGE; ++i) { mbar_init(smem_u32(&bars->full[i]), 1); mbar_init(smem_u32(&bars->empty[i]), 1); }
    for (int i = 0; i < NBUF; ++i) { mbar_init(smem_u32(&bars->tmem_full[i]), 1); mbar_init(smem_u32(&bars->tmem_empty[i]), 2 * NEPI); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) tmem_alloc_pair(smem_u32(&bars->tmem_base), 512);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();          // the peer's barriers exist before anything is signalled on them
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  if (threadIdx.x == 0) TL(2);

  if (warp == 0) {
    // ================================ TMA producer (both CTAs) ================================
    if (elect_one()) {
      int it = 0, used = 0;
      for (int w = pair0; w < P.nwork; w += npairs) {
        const PairWork k = pair_work<R>(P, w);
        if (k.nblk == 0) continue;
        const int y0 = k.y0p + (int)rank * TILE_R;
        const int ya = y0 < k.Hc ? y0 : 0;   // a phantom second tile (odd number of row tiles) loads valid rows, unused
        for (int blk = 0; blk < k.nblk; ++blk) {
          const int srow = k.s_first + blk * 4 + 2 * (int)rank;   // this CTA's two b-rows of the block
          for (int kc = 0; kc < nkc; ++kc, ++it) {
            if (blk == 0) {   // the a tile goes out chunk by chunk, interleaved with block 0 which consumes it in this order
              if (used > 0) mbar_wait(smem_u32(&bars->a_free[kc]), (used - 1) & 1);   // the previous tile is done with it
              if (rank == 0) mbar_expect_tx(smem_u32(&bars->a_chunk[kc]), 2u * CHUNK_BYTES);
              tma_load_4d_pair(smem_u32(sA + (size_t)kc * CHUNK_BYTES), &P.a[k.cls],
                               mapa_u32(smem_u32(&bars->a_chunk[kc]), 0), kc * KCH, x0, ya, k.n);
            }
            const int stage = it % PNSTAGE;
            mbar_wait(smem_u32(&bars->empty[stage]), ((it / PNSTAGE) & 1) ^ 1);
            if (rank == 0) mbar_expect_tx(smem_u32(&bars->full[stage]), 2u * PSTAGE_BYTES);
            tma_load_4d_pair(smem_u32(sB + (size_t)stage * PSTAGE_BYTES), &P.b2[k.cls],
                             mapa_u32(smem_u32(&bars->full[stage]), 0), kc * KCH, x0, srow, k.n);
          }
        }
        ++used;
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (leader only) ================================
    if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = idesc_tf32(256, BLK_N);
      int it = 0, gb = 0, used = 0;
      for (int w = pair0; w < P.nwork; w += npairs) {
        const PairWork k = pair_work<R>(P, w);
        if (k.nblk == 0) continue;
        for (int blk = 0; blk < k.nblk; ++blk, ++gb) {
          const int buf = gb % NBUF;
          mbar_wait(smem_u32(&bars->tmem_empty[buf]), ((gb / NBUF) & 1) ^ 1);
          tc_fence_after();
          if (used == 0) TL(4 + blk);
          const uint32_t d_tmem = tmem_base + (uint32_t)buf * BLK_N;
          for (int kc = 0; kc < nkc; ++kc, ++it) {
            const int stage = it % PNSTAGE;
            if (blk == 0) mbar_wait(smem_u32(&bars->a_chunk[kc]), used & 1);
            mbar_wait(smem_u32(&bars->full[stage]), (it / PNSTAGE) & 1);
            tc_fence_after();
            if (kc == 0 && used == 0) TL(12 + blk);
            const uint32_t a_addr = smem_u32(sA + (size_t)kc * CHUNK_BYTES);
            const uint32_t b_addr = smem_u32(sB + (size_t)stage * PSTAGE_BYTES);
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
              umma_tf32_pair(d_tmem, desc_sw128_kmajor(a_addr + k4 * 32), desc_sw128_kmajor(b_addr + k4 * 32), idesc,
                             (uint32_t)((kc | k4) != 0));
            umma_commit_pair(smem_u32(&bars->empty[stage]), 3);   // frees the stage in both CTAs
            if (blk == k.nblk - 1) umma_commit_pair(smem_u32(&bars->a_free[kc]), 3);   // a chunk kc may be replaced
          }
          umma_commit_pair(smem_u32(&bars->tmem_full[buf]), 3);   // accumulator block complete in both CTAs
          if (used == 0) TL(20 + blk);
        }
        ++used;
      }
    }
  } else {
    // ================================ epilogue (warps 2..9, both CTAs) ================================
    const int ew = warp - 2;
    const int q = warp & 3;              // TMEM lane quadrant this warp may read = a-row of the tile
    const int half = ew >> 2;            // the two warps of a quadrant take alternate b-rows
    float* stg = sStage + ew * STG_FLOATS;
    const float inv = 1.0f / P.sumelems;
    const float slope = P.slope;
    const int pixstride = P.s2 * P.opitch;
    const uint32_t tmem_empty0 = mapa_u32(smem_u32(&bars->tmem_empty[0]), 0);

    // per-lane plan for writing one displacement row (32 pixels x GW values) coalesced; see corr_tc_fwd_kernel.
    // rd = staging word to read (the never-written pad word 32, zeroed below, for displacements that leave the image
    // sideways), go = byte offset of the output element from the row's first element, cq = its pixel column.
    int rd[GW];
    uint32_t go[GW];
#pragma unroll
    for (int tt = 0; tt < GW; ++tt) {
      const int e = tt * 32 + lane;
      const int cc = e / GW, o = e - cc * GW;
      const int xb = cc + o - R;
      const bool xok = (xb >= 0) && (xb < 32);
      rd[tt] = xok ? xb * 33 + cc : 32;
      go[tt] = (uint32_t)(cc * pixstride + o) * 4u;
    }
    if (lane == 0) stg[32] = 0.f;
    __syncwarp();

    int gb = 0;
    for (int w = pair0; w < P.nwork; w += npairs) {
      const PairWork k = pair_work<R>(P, w);
      if (!k.valid) continue;
      const int y = k.y0p + (int)rank * TILE_R + q;
      const bool row_ok = y < k.Hc;
      const int Y = y * P.s2 + k.cy;
      const int ncol = min(TILE_C, k.Wc - x0);
      float* orow0 = P.out + (((size_t)k.n * P.H + (row_ok ? Y : 0)) * P.W + (size_t)(x0 * P.s2 + k.cx)) * P.opitch;
      uint32_t smask = 0;                // stores of pixel columns that exist (partial tiles only)
      if (ncol != TILE_C) {
#pragma unroll
        for (int tt = 0; tt < GW; ++tt) smask |= (((tt * 32 + lane) / GW < ncol) ? 1u : 0u) << tt;
      }

      if (row_ok && k.first_part) {   // displacement rows that point outside the image: zeros, never computed
        const int yb = y + P.boff;
        const int dlo = min(GW, max(0, R - yb)) * GW;
        const int dhi = max(0, min(GW, k.Hcb - yb + R)) * GW;
        for (int cc = half; cc < ncol; cc += 2) {
          float* po = orow0 + (size_t)cc * pixstride;
          for (int e = lane; e < dlo; e += 32) po[e] = 0.f;
          for (int e = dhi + lane; e < P.D; e += 32) po[e] = 0.f;
        }
      }

      for (int blk = 0; blk < k.nblk; ++blk, ++gb) {
        const int buf = gb % NBUF;
        mbar_wait(smem_u32(&bars->tmem_full[buf]), (gb / NBUF) & 1);
        tc_fence_after();
        if (ew == 0 && lane == 0 && w == pair0) TL(28 + blk);
#pragma unroll
        for (int h2 = 0; h2 < 2; ++h2) {
          const int br = half + 2 * h2;
          const int s = k.s_first + blk * 4 + br;
          const int p = s - (y + P.boff);
          if (row_ok && p >= -R && p <= R && s < k.Hcb) {   // warp-uniform
            float v[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BLK_N + br * 32), v);
#pragma unroll
            for (int j = 0; j < 32; ++j) stg[j * 33 + lane] = v[j];
            __syncwarp();
            char* orow = reinterpret_cast<char*>(orow0 + (p + R) * GW);
            if (ncol == TILE_C) {   // full tile (warp-uniform): straight-line load, scale, store
#pragma unroll
              for (int tt = 0; tt < GW; ++tt)
                *reinterpret_cast<float*>(orow + go[tt]) = leaky(stg[rd[tt]] * inv, slope);
            } else {
#pragma unroll
              for (int tt = 0; tt < GW; ++tt)
                if ((smask >> tt) & 1u) *reinterpret_cast<float*>(orow + go[tt]) = leaky(stg[rd[tt]] * inv, slope);
            }
            __syncwarp();
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(tmem_empty0 + (uint32_t)buf * 8u);
        if (ew == 0 && lane == 0 && w == pair0) TL(36 + blk);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) TL(44);
  cluster_sync_all();          // neither CTA may go while the other can still touch its shared memory / barriers
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 512);
  }
  if (threadIdx.x == 32) {
    TL(45); TLG(46);
#ifdef FN2_TIMELINE
    unsigned sm_; asm volatile("mov.u32 %0, %%smid;" : "=r"(sm_));
    if (blockIdx.x < 1024) { g_tl[blockIdx.x * 64 + 47] = sm_; g_tl[blockIdx.x * 64 + 48] = pair_work<R>(P, pair0).nblk; }
#endif
  }
}

size_t fwd_pair_smem_bytes(int nkc) {
  return 1024 + (size_t)nkc * CHUNK_BYTES + (size_t)PNSTAGE * PSTAGE_BYTES + (size_t)NEPI * STG_FLOATS * sizeof(float) +
         sizeof(PairBarriers);
}

size_t fwd_smem_bytes(int nkc) {
  return 1024 + (size_t)nkc * CHUNK_BYTES + (size_t)NSTAGE * CHUNK_BYTES + (size_t)NEPI * STG_FLOATS * sizeof(float) +
         sizeof(Barriers);
}

}  // namespace

bool corr_tc_supported(const CorrGeom& g, int precision) {
  if (precision != FN2_PREC_TF32 && precision != FN2_PREC_TF32X3) return false;
  if (g.k != 1 || g.s1 != 1 || g.pad != g.md) return false;
  if (g.s2 != 1 && g.s2 != 2) return false;
  if (g.r != 10) return false;                       // window radius instantiated below
  if (g.C % KCH != 0 || g.C / KCH > MAX_KC) return false;
  if (g.bTop % g.s2 != 0) return false;              // a row halo must keep the parity classes aligned
  return true;
}

size_t corr_tc_workspace_bytes(const CorrGeom&, int, int) { return 0; }

int corr_tc_forward(const CorrGeom& g, const float* a, const float* b, float* out, int precision, void*, size_t,
                    cudaStream_t st) {
  if (!corr_tc_supported(g, precision)) return FN2_E_UNSUPPORTED_PRECISION;
  if (precision == FN2_PREC_TF32X3) return corr_tc_x3_forward(g, a, b, out, st);   // corr_tc_x3.cu
  TcParams P;
  P.out = out;
  P.N = g.N; P.H = g.H; P.W = g.W; P.C = g.C; P.D = g.D; P.s2 = g.s2; P.ncls = g.s2 * g.s2;
  P.boff = g.bTop / g.s2;
  for (int c = 0; c < 2; ++c) {
    P.Hc[c] = c < g.s2 ? (g.H - c + g.s2 - 1) / g.s2 : 0;
    P.Wc[c] = c < g.s2 ? (g.W - c + g.s2 - 1) / g.s2 : 0;
    P.Hcb[c] = c < g.s2 ? (g.bH - c + g.s2 - 1) / g.s2 : 0;
  }
  P.TY = (P.Hc[0] + TILE_R - 1) / TILE_R;
  P.TX = (P.Wc[0] + TILE_C - 1) / TILE_C;
  P.TYP = (P.TY + 1) / 2;
  P.nkc = g.C / KCH;
  P.sumelems = (float)g.C;
  P.opitch = g.oPitch;
  P.slope = g.slope;
  for (int cls = 0; cls < P.ncls; ++cls) {
    const int cy = cls / g.s2, cx = cls % g.s2;
    if (P.Hc[cy] < 1 || P.Wc[cx] < 1) {  // empty class (H or W == 1 with s2 == 2): map class 0 instead, never used
      P.a[cls] = P.a[0]; P.b[cls] = P.b[0]; P.b2[cls] = P.b2[0];
      continue;
    }
    if (!make_class_map(&P.a[cls], a, g.N, g.H, g.W, g.C, g.s2, cy, cx, P.Hc[cy], P.Wc[cx], TILE_C, TILE_R) ||
        !make_class_map(&P.b[cls], b, g.N, g.bH, g.W, g.C, g.s2, cy, cx, P.Hcb[cy], P.Wc[cx], TILE_C, 4) ||
        !make_class_map(&P.b2[cls], b, g.N, g.bH, g.W, g.C, g.s2, cy, cx, P.Hcb[cy], P.Wc[cx], TILE_C, 2))
      return FN2_E_CUDA;
  }
  for (int cls = P.ncls; cls < 4; ++cls) { P.a[cls] = P.a[0]; P.b[cls] = P.b[0]; P.b2[cls] = P.b2[0]; }
  const size_t smem = fwd_smem_bytes(P.nkc);
  static PerDeviceOnce once;   // function attributes are per device
  const cudaError_t attr_err = per_device_once(once, [] {
    cudaError_t e = cudaFuncSetAttribute(corr_tc_fwd_kernel<10, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)fwd_smem_bytes(MAX_KC));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(corr_tc_fwd_kernel<10, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)fwd_smem_bytes(MAX_KC));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(corr_tc_fwd_pair_kernel<10>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)fwd_pair_smem_bytes(MAX_KC));
    return e;
  });
  if (attr_err != cudaSuccess) { note_cuda_error(attr_err); (void)cudaGetLastError(); return FN2_E_CUDA; }
  const long long grid = (long long)g.N * P.ncls * P.TY * P.TX;
  if (grid > 0x7fffffffLL) return FN2_E_BAD_SHAPE;
  const EnvSwitches& sw = env_switches();
  if (P.TX == 1 && !sw.no_pair) {   // narrow images: CTA pairs share every b block
    // Tail split: with `slots` CTA pairs resident at a time, the last npairs % slots pair tiles would run alone in
    // a final wave; cut each into pieces of g blocks (smallest g whose pieces all fit in that wave).
    const int slots = std::max(device_sm_count() / 2, 1);
    const long long npairs = (long long)g.N * P.ncls * P.TYP;
    const int rows_max = std::min(P.Hcb[0], 2 * TILE_R + 2 * 10);
    const int maxblk = (rows_max + 3) / 4;
    const int tail_tiles = (int)(npairs % slots);
    P.n_full = (int)npairs; P.split_units = 1; P.split_g = maxblk;
    if (npairs > slots && tail_tiles > 0 && !sw.no_split) {
      for (int gsz = 1; gsz < maxblk; ++gsz) {
        const int units = (maxblk + gsz - 1) / gsz;
        if ((long long)tail_tiles * units <= slots) {
          P.n_full = (int)(npairs - tail_tiles); P.split_units = units; P.split_g = gsz;
          break;
        }
      }
    }
    const long long nwork = (long long)P.n_full + (npairs - P.n_full) * P.split_units;
    if (nwork > 0x3fffffffLL) return FN2_E_BAD_SHAPE;
    P.nwork = (int)nwork;
    // persistent: one pair per pair of SMs strides over the work list (FN2_NO_PERSIST: one pair per unit of work)
    const long long pgrid = 2 * (sw.no_persist ? nwork : std::min<long long>(nwork, slots));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)pgrid);
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = fwd_pair_smem_bytes(P.nkc);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, corr_tc_fwd_pair_kernel<10>, P);
    if (e != cudaSuccess) { note_cuda_error(e); (void)cudaGetLastError(); return FN2_E_CUDA; }
    return after_launch();
  }
  if (P.TX > 1)
    corr_tc_fwd_kernel<10, true><<<(unsigned)grid, NTHREADS, smem, st>>>(P);
  else
    corr_tc_fwd_kernel<10, false><<<(unsigned)grid, NTHREADS, smem, st>>>(P);
  return after_launch();
}

#ifdef FN2_TIMELINE
extern "C" __attribute__((visibility("default"))) int fn2_debug_timeline(long long* host_out) {
  return (int)cudaMemcpyFromSymbol(host_out, g_tl, sizeof(long long) * 1024 * 64);
}
#endif

int corr_tc_backward(const CorrGeom& g, const float* gout, const float* a, const float* b, float* ga, float* gb,
                     int precision, void*, size_t, cudaStream_t st) {
  if (!corr_tc_supported(g, precision)) return FN2_E_UNSUPPORTED_PRECISION;
  if (corr_tc_bwd_supported(g, precision)) return corr_tc_backward_launch(g, gout, a, b, ga, gb, precision, st);
  // shapes the tensor-core backward does not cover: the fp32 CUDA-core path keeps the op differentiable
  return corr_generic_backward(g, gout, a, b, ga, gb, st);
}

}  // namespace fn2
