// Microbenchmark (developer tool, not part of the library): ways to accumulate the four-tap image-gradient scatter of
// flow_warp backward (flow_warp_grad.cu.cc:43-53) for a 32x64 tile of source pixels, C = 3, flows ~ N(0, 8 px).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mb_scatter mb_scatter.cu && ./mb_scatter
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include <stdint.h>

constexpr int TH = 32, TW = 64, NT = 512, PXT = 4, C = 3;
constexpr int MY = 22, MX = 24, ROWS = TH + 2 * MY + 1, COLS = 120;
constexpr int WF = ROWS * COLS * C;   // floats in the window

struct Taps { bool in; int L, T, R, B; float a, b; };
__device__ __forceinline__ Taps make_taps(float fx, float fy, int x, int y, int W, int H) {
  Taps t; const float x2 = x + fx, y2 = y + fy;
  t.in = (x2 >= 0.f) && (y2 >= 0.f) && (x2 < (float)W) && (y2 < (float)H);
  t.L = t.in ? (int)x2 : 0; t.T = t.in ? (int)y2 : 0;
  t.R = min(t.L + 1, W - 1); t.B = min(t.T + 1, H - 1);
  t.a = x2 - t.L; t.b = y2 - t.T; return t;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// MODE 0: fp32 CAS atomics in smem; 1: one int32 native atomic per value (timing only); 2: two int32 (hi/lo) per value
template <int MODE>
__global__ void __launch_bounds__(NT) k_smem(const float* __restrict__ flow, const float* __restrict__ gout,
                                             float* __restrict__ gimage, int H, int W, int tilesX, int tilesY) {
  extern __shared__ __align__(128) float win[];
  int b = blockIdx.x; const int txi = b % tilesX; b /= tilesX; const int tyi = b % tilesY; const int n = b / tilesY;
  const int ty0 = tyi * TH, tx0 = txi * TW, wy0 = ty0 - MY, wx0 = tx0 - MX;
  const int y = ty0 + threadIdx.x / (TW / PXT), x = tx0 + (threadIdx.x % (TW / PXT)) * PXT;
  const size_t pix = ((size_t)n * H + y) * W + x;
  const float4 f01 = __ldg((const float4*)(flow + pix * 2)), f23 = __ldg((const float4*)(flow + pix * 2) + 1);
  const float fx[4] = {f01.x, f01.z, f23.x, f23.z}, fy[4] = {f01.y, f01.w, f23.y, f23.w};
  float gv[12];
  for (int v = 0; v < 3; ++v) { const float4 q = __ldg((const float4*)(gout + pix * 3) + v); gv[4*v]=q.x; gv[4*v+1]=q.y; gv[4*v+2]=q.z; gv[4*v+3]=q.w; }
  constexpr int NW = MODE == 2 ? 2 * WF : WF;
  for (int i = threadIdx.x; i < NW / 4; i += NT) ((float4*)win)[i] = make_float4(0, 0, 0, 0);
  __syncthreads();
  const float s1 = 32768.f, q1 = 1.f / 32768.f, s2 = 32768.f * 65536.f;
  float* gimg = gimage + (size_t)n * H * W * C;
#pragma unroll
  for (int i = 0; i < PXT; ++i) {
    const Taps tp = make_taps(fx[i], fy[i], x + i, y, W, H);
    if (!tp.in) continue;
    const float cf[4] = {(1 - tp.a) * (1 - tp.b), tp.a * (1 - tp.b), (1 - tp.a) * tp.b, tp.a * tp.b};
    const bool inwin = tp.T >= wy0 && tp.B < wy0 + ROWS && tp.L >= wx0 && tp.R < wx0 + COLS;
    if (inwin) {
      const int oL = ((tp.T - wy0) * COLS + (tp.L - wx0)) * C, oR = ((tp.T - wy0) * COLS + (tp.R - wx0)) * C;
      const int dB = (tp.B - tp.T) * COLS * C;
      const int offs[4] = {oL, oR, oL + dB, oR + dB};
#pragma unroll
      for (int t = 0; t < 4; ++t)
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const float val = gv[i * C + c] * cf[t];
          if (MODE == 0) atomicAdd(win + offs[t] + c, val);
          else if (MODE == 3) {   // optimistic CAS: guess that the word is still zero, no load first
            unsigned* a = (unsigned*)(win + offs[t] + c);
            unsigned assumed = 0u, old;
            while ((old = atomicCAS(a, assumed, __float_as_uint(__uint_as_float(assumed) + val))) != assumed) assumed = old;
          }
          else if (MODE == 1) atomicAdd((int*)win + offs[t] + c, __float2int_rn(val * 1048576.f));
          else {
            const float hf = rintf(val * s1);
            const float r = fmaf(-hf, q1, val);
            atomicAdd((int*)win + 2 * (offs[t] + c), (int)hf);
            atomicAdd((int*)win + 2 * (offs[t] + c) + 1, __float2int_rn(r * s2));
          }
        }
    } else {
      float* pT = gimg + ((size_t)tp.T * W + tp.L) * C;
      const size_t dR = (size_t)(tp.R - tp.L) * C, dB = (size_t)(tp.B - tp.T) * W * C;
      for (int c = 0; c < C; ++c) {
        const float gc = gv[i * C + c];
        atomicAdd(pT + c, gc * cf[0]); atomicAdd(pT + dR + c, gc * cf[1]);
        atomicAdd(pT + dB + c, gc * cf[2]); atomicAdd(pT + dB + dR + c, gc * cf[3]);
      }
    }
  }
  __syncthreads();
  if (MODE == 1) { for (int i = threadIdx.x; i < WF; i += NT) win[i] = (float)((int*)win)[i] * (1.f / 1048576.f); }
  if (MODE == 2) {   // in-place compaction is unsafe in general; for timing: convert into the first half row by row
    for (int r = 0; r < ROWS; ++r) {
      float val = 0.f;
      const int i = threadIdx.x;
      if (i < COLS * C) { const int2 v = ((int2*)win)[r * COLS * C + i]; val = (float)v.x * q1 + (float)v.y * (1.f / s2); }
      __syncthreads();
      if (i < COLS * C) win[r * COLS * C + i] = val;
      __syncthreads();
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < ROWS) {
    const int r = threadIdx.x, yy = wy0 + r;
    const int fc0 = wx0 * C, lo = max(fc0, 0), hi = min(fc0 + COLS * C, W * C);
    if (yy >= 0 && yy < H && hi > lo) {
      float* dst = gimage + ((size_t)n * H + yy) * W * C + lo;
      const uint32_t src = smem_u32(win + r * COLS * C + (lo - fc0));
      asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"((uint32_t)(hi - lo) * 4u) : "memory");
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
}

// MODE 0: scalar red (the reference's way); 1: 8-byte pairs where aligned; 2: zero-padded aligned 16-byte groups
template <int MODE>
__global__ void __launch_bounds__(NT) k_gmem(const float* __restrict__ flow, const float* __restrict__ gout,
                                             float* __restrict__ gimage, int H, int W, int tilesX, int tilesY) {
  int b = blockIdx.x; const int txi = b % tilesX; b /= tilesX; const int tyi = b % tilesY; const int n = b / tilesY;
  const int y = tyi * TH + threadIdx.x / (TW / PXT), x = txi * TW + (threadIdx.x % (TW / PXT)) * PXT;
  const size_t pix = ((size_t)n * H + y) * W + x;
  const float4 f01 = __ldg((const float4*)(flow + pix * 2)), f23 = __ldg((const float4*)(flow + pix * 2) + 1);
  const float fx[4] = {f01.x, f01.z, f23.x, f23.z}, fy[4] = {f01.y, f01.w, f23.y, f23.w};
  float gv[12];
  for (int v = 0; v < 3; ++v) { const float4 q = __ldg((const float4*)(gout + pix * 3) + v); gv[4*v]=q.x; gv[4*v+1]=q.y; gv[4*v+2]=q.z; gv[4*v+3]=q.w; }
  float* gimg = gimage + (size_t)n * H * W * C;
#pragma unroll
  for (int i = 0; i < PXT; ++i) {
    const Taps tp = make_taps(fx[i], fy[i], x + i, y, W, H);
    if (!tp.in) continue;
    const float cTL = (1 - tp.a) * (1 - tp.b), cTR = tp.a * (1 - tp.b), cBL = (1 - tp.a) * tp.b, cBR = tp.a * tp.b;
    const bool edge = (tp.R == tp.L) || (tp.B == tp.T);
    if (MODE == 0 || edge) {
      float* pT = gimg + ((size_t)tp.T * W + tp.L) * C;
      const size_t dR = (size_t)(tp.R - tp.L) * C, dB = (size_t)(tp.B - tp.T) * W * C;
      for (int c = 0; c < C; ++c) {
        const float gc = gv[i * C + c];
        atomicAdd(pT + c, gc * cTL); atomicAdd(pT + dR + c, gc * cTR);
        atomicAdd(pT + dB + c, gc * cBL); atomicAdd(pT + dB + dR + c, gc * cBR);
      }
    } else {
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) {
        const float cl = rr ? cBL : cTL, cr = rr ? cBR : cTR;
        float v6[6];
        for (int c = 0; c < C; ++c) { v6[c] = gv[i * C + c] * cl; v6[3 + c] = gv[i * C + c] * cr; }
        const size_t o = ((size_t)(rr ? tp.B : tp.T) * W + tp.L) * C;   // float offset inside the image (image base 16B aligned, H*W*C*4 % 16 == 0)
        float* p = gimg + o;
        if (MODE == 1) {
          if ((o & 1) == 0) {
            for (int k = 0; k < 3; ++k) atomicAdd((float2*)(p + 2 * k), make_float2(v6[2 * k], v6[2 * k + 1]));
          } else {
            atomicAdd(p, v6[0]);
            atomicAdd((float2*)(p + 1), make_float2(v6[1], v6[2]));
            atomicAdd((float2*)(p + 3), make_float2(v6[3], v6[4]));
            atomicAdd(p + 5, v6[5]);
          }
        } else {
          const int m = (int)(o & 3);
          float e[12];
          for (int k = 0; k < 12; ++k) e[k] = 0.f;
          for (int k = 0; k < 6; ++k) {   // e[m + k] = v6[k] without dynamic register indexing
            for (int mm = 0; mm < 4; ++mm) if (m == mm) e[mm + k] = v6[k];
          }
          float* q = p - m;
          atomicAdd((float4*)q, make_float4(e[0], e[1], e[2], e[3]));
          atomicAdd((float4*)(q + 4), make_float4(e[4], e[5], e[6], e[7]));
          if (m == 3) atomicAdd((float4*)(q + 8), make_float4(e[8], e[9], e[10], e[11]));
        }
      }
    }
  }
}


// fp32 CAS window with margins (MYt, MXt); taps outside the window go to global memory as 8-byte reds where aligned.
// SPLIT: additionally send the BOTTOM tap row of every pixel to global memory (both atomic units busy).
template <int MYt, int MXt, int SPLIT>
__global__ void __launch_bounds__(NT) k_win(const float* __restrict__ flow, const float* __restrict__ gout,
                                            float* __restrict__ gimage, int H, int W, int tilesX, int tilesY) {
  constexpr int ROWSt = TH + 2 * MYt + 1, COLSt = (TW + 2 * MXt + 1 + 3) / 4 * 4, WFt = ROWSt * COLSt * C;
  extern __shared__ __align__(128) float win[];
  int b = blockIdx.x; const int txi = b % tilesX; b /= tilesX; const int tyi = b % tilesY; const int n = b / tilesY;
  const int ty0 = tyi * TH, tx0 = txi * TW, wy0 = ty0 - MYt, wx0 = tx0 - MXt;
  const int y = ty0 + threadIdx.x / (TW / PXT), x = tx0 + (threadIdx.x % (TW / PXT)) * PXT;
  const size_t pix = ((size_t)n * H + y) * W + x;
  const float4 f01 = __ldg((const float4*)(flow + pix * 2)), f23 = __ldg((const float4*)(flow + pix * 2) + 1);
  const float fx[4] = {f01.x, f01.z, f23.x, f23.z}, fy[4] = {f01.y, f01.w, f23.y, f23.w};
  float gv[12];
  for (int v = 0; v < 3; ++v) { const float4 q = __ldg((const float4*)(gout + pix * 3) + v); gv[4*v]=q.x; gv[4*v+1]=q.y; gv[4*v+2]=q.z; gv[4*v+3]=q.w; }
  for (int i = threadIdx.x; i < WFt / 4; i += NT) ((float4*)win)[i] = make_float4(0, 0, 0, 0);
  __syncthreads();
  float* gimg = gimage + (size_t)n * H * W * C;
#pragma unroll
  for (int i = 0; i < PXT; ++i) {
    const Taps tp = make_taps(fx[i], fy[i], x + i, y, W, H);
    if (!tp.in) continue;
    if (SPLIT == 2) { if (tp.a == 123.f) win[i] = tp.b; continue; }
    const float cf[4] = {(1 - tp.a) * (1 - tp.b), tp.a * (1 - tp.b), (1 - tp.a) * tp.b, tp.a * tp.b};
    const bool inwin = tp.T >= wy0 && tp.B < wy0 + ROWSt && tp.L >= wx0 && tp.R < wx0 + COLSt;
    const bool edge = (tp.R == tp.L) || (tp.B == tp.T);
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {
      const float cl = cf[2 * rr], cr = cf[2 * rr + 1];
      const int yy = rr ? tp.B : tp.T;
      if (inwin && !(SPLIT == 1 && rr == 1)) {
        float* pL = win + ((yy - wy0) * COLSt + (tp.L - wx0)) * C;
        float* pR = win + ((yy - wy0) * COLSt + (tp.R - wx0)) * C;
#pragma unroll
        for (int c = 0; c < C; ++c) { atomicAdd(pL + c, gv[i * C + c] * cl); atomicAdd(pR + c, gv[i * C + c] * cr); }
      } else if (edge) {
        float* pL = gimg + ((size_t)yy * W + tp.L) * C;
        float* pR = gimg + ((size_t)yy * W + tp.R) * C;
        for (int c = 0; c < C; ++c) { atomicAdd(pL + c, gv[i * C + c] * cl); atomicAdd(pR + c, gv[i * C + c] * cr); }
      } else {
        float v6[6];
        for (int c = 0; c < C; ++c) { v6[c] = gv[i * C + c] * cl; v6[3 + c] = gv[i * C + c] * cr; }
        const size_t o = ((size_t)yy * W + tp.L) * C;
        float* p = gimg + o;
        if ((o & 1) == 0) {
          for (int k = 0; k < 3; ++k) atomicAdd((float2*)(p + 2 * k), make_float2(v6[2 * k], v6[2 * k + 1]));
        } else {
          atomicAdd(p, v6[0]);
          atomicAdd((float2*)(p + 1), make_float2(v6[1], v6[2]));
          atomicAdd((float2*)(p + 3), make_float2(v6[3], v6[4]));
          atomicAdd(p + 5, v6[5]);
        }
      }
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < ROWSt) {
    const int r = threadIdx.x, yy = wy0 + r;
    const int fc0 = wx0 * C, lo = max(fc0, 0), hi = min(fc0 + COLSt * C, W * C);
    if (yy >= 0 && yy < H && hi > lo) {
      float* dst = gimage + ((size_t)n * H + yy) * W * C + lo;
      const uint32_t src = smem_u32(win + r * COLSt * C + (lo - fc0));
      asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"((uint32_t)(hi - lo) * 4u) : "memory");
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
}
template <int MYt, int MXt, int SPLIT>
float run_win(const float* flow, const float* gout, float* gimg, int N, int H, int W, size_t npx, cudaEvent_t e0, cudaEvent_t e1) {
  constexpr int ROWSt = TH + 2 * MYt + 1, COLSt = (TW + 2 * MXt + 1 + 3) / 4 * 4, WFt = ROWSt * COLSt * C;
  cudaFuncSetAttribute(k_win<MYt, MXt, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, WFt * 4 + 128);
  float best = 1e9f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaMemsetAsync(gimg, 0, npx * 12);
    cudaEventRecord(e0);
    k_win<MYt, MXt, SPLIT><<<N * (W / TW) * (H / TH), NT, WFt * 4 + 128>>>(flow, gout, gimg, H, W, W / TW, H / TH);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); best = fminf(best, ms);
  }
  return best;
}

static float gauss() { float u = (rand() + 1.f) / (RAND_MAX + 2.f), v = (rand() + 1.f) / (RAND_MAX + 2.f); return sqrtf(-2 * logf(u)) * cosf(6.2831853f * v); }

int main(int argc, char** argv) {
  const int N = 32, H = 384, W = 512;
  const float sigma = argc > 1 ? atof(argv[1]) : 8.f;
  const size_t npx = (size_t)N * H * W;
  std::vector<float> hf(npx * 2), hg(npx * 3);
  srand(1);
  for (auto& v : hf) v = gauss() * sigma;
  for (auto& v : hg) v = gauss();
  float *flow, *gout, *gimg, *ref;
  cudaMalloc(&flow, npx * 8); cudaMalloc(&gout, npx * 12); cudaMalloc(&gimg, npx * 12); cudaMalloc(&ref, npx * 12);
  cudaMemcpy(flow, hf.data(), npx * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(gout, hg.data(), npx * 12, cudaMemcpyHostToDevice);
  const int tilesX = W / TW, tilesY = H / TH, blocks = N * tilesX * tilesY;
  cudaFuncSetAttribute(k_smem<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, WF * 4 + 128);
  cudaFuncSetAttribute(k_smem<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, WF * 4 + 128);
  cudaFuncSetAttribute(k_smem<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, WF * 4 + 128);
  cudaFuncSetAttribute(k_smem<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, WF * 8 + 128);
  std::vector<float> hr(npx * 3), ho(npx * 3);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int variant = 0; variant < 7; ++variant) {
    float best = 1e9f;
    for (int rep = 0; rep < 5; ++rep) {
      cudaMemsetAsync(gimg, 0, npx * 12);
      cudaEventRecord(e0);
      switch (variant) {
        case 0: k_smem<0><<<blocks, NT, WF * 4 + 128>>>(flow, gout, gimg, H, W, tilesX, tilesY); break;
        case 1: k_smem<1><<<blocks, NT, WF * 4 + 128>>>(flow, gout, gimg, H, W, tilesX, tilesY); break;
        case 2: k_smem<2><<<blocks, NT, WF * 8 + 128>>>(flow, gout, gimg, H, W, tilesX, tilesY); break;
        case 3: k_gmem<0><<<blocks, NT>>>(flow, gout, gimg, H, W, tilesX, tilesY); break;
        case 4: k_gmem<1><<<blocks, NT>>>(flow, gout, gimg, H, W, tilesX, tilesY); break;
        case 5: k_gmem<2><<<blocks, NT>>>(flow, gout, gimg, H, W, tilesX, tilesY); break;
        case 6: k_smem<3><<<blocks, NT, WF * 4 + 128>>>(flow, gout, gimg, H, W, tilesX, tilesY); break;
      }
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1); best = fminf(best, ms);
    }
    cudaError_t err = cudaGetLastError();
    cudaMemcpy(ho.data(), gimg, npx * 12, cudaMemcpyDeviceToHost);
    if (variant == 0) hr = ho;
    double md = 0; for (size_t i = 0; i < ho.size(); i += 7) md = fmax(md, fabs((double)ho[i] - hr[i]));
    const char* names[] = {"smem fp32 CAS", "smem int32 x1 (timing only)", "smem int32 hi/lo", "gmem red scalar", "gmem red x2 where aligned", "gmem red x4 padded", "smem optimistic CAS (guess 0)"};
    printf("sigma %.1f  %-30s %.4f ms  maxdiff vs variant0 %.3g  %s\n", sigma, names[variant], best, md, cudaGetErrorString(err));
  }

  {
    struct { const char* name; float ms; } r[] = {
      {"window 22x24", run_win<22, 24, 0>(flow, gout, gimg, N, H, W, npx, e0, e1)},
      {"window 16x16", run_win<16, 16, 0>(flow, gout, gimg, N, H, W, npx, e0, e1)},
      {"window 12x12", run_win<12, 12, 0>(flow, gout, gimg, N, H, W, npx, e0, e1)},
      {"window 8x8", run_win<8, 8, 0>(flow, gout, gimg, N, H, W, npx, e0, e1)},
      {"window 4x4", run_win<4, 4, 0>(flow, gout, gimg, N, H, W, npx, e0, e1)},
      {"window 22x24 + bottom row to gmem", run_win<22, 24, 1>(flow, gout, gimg, N, H, W, npx, e0, e1)},
      {"window 12x12 + bottom row to gmem", run_win<12, 12, 1>(flow, gout, gimg, N, H, W, npx, e0, e1)},
      {"window 22x24, NO atomics (zero+flush)", run_win<22, 24, 2>(flow, gout, gimg, N, H, W, npx, e0, e1)},
      {"window 4x4, NO atomics (zero+flush)", run_win<4, 4, 2>(flow, gout, gimg, N, H, W, npx, e0, e1)},
    };
    for (auto& q : r) printf("sigma %.1f  %-36s %.4f ms  %s\n", sigma, q.name, q.ms, cudaGetErrorString(cudaGetLastError()));
    cudaMemcpy(ho.data(), gimg, npx * 12, cudaMemcpyDeviceToHost);
    double md = 0; for (size_t i = 0; i < ho.size(); i += 7) md = fmax(md, fabs((double)ho[i] - hr[i]));
    printf("last variant maxdiff vs variant0 %.3g\n", md);
  }
  return 0;
}
