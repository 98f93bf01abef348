// cube_f32_kernel.cu -- K4: the reference's float CNN (Cube.AI build of kws_conv) in fp32 FFMA.
//
// Reference: topology audio/edison/train/kws_keras.py:371-397; generated network firmware/src/ai/cube/kws/kws.c:306-420
// (conv2d_nl_pool x2, conv2d x2, dense, forward_sm), tensors :213-291, weight offsets :488-516; entry ai_kws_run via
// cubeNetRun firmware/src/ai/ai.c:354-378; input = MFCC values cast to float, unscaled (app.c:679-683).
// The arithmetic of ST's closed runtime is unknown beyond "float32 MACs": parity is tolerance-based against the fp64
// restatement in the test tree (cube_float.py) and UNPINNED at the keras boundary.
//
// A persistent CTA per SM keeps a tile of 8 clips in shared memory through all layers (1,612 B in, 40 + 40 B out per clip) and
// ALL conv weights beside it: w1/w2 (20 KB) stay for the CTA's life, w3 and w4 (73.7 KB each) take turns in one buffer, refilled
// by cp.async -- w3 while conv1/conv2 of the same tile run, w4 between conv3 and conv4 (~2 % of a tile, not overlapped) -- and
// the next tile's input is prefetched into its staging buffer as soon as conv1 has consumed the current one.  (Weights read
// through L1 instead missed 7 % of the time -- 172 KB of weights against the ~80 KB of L1 left beside the tile -- and those
// L2 round trips were the top stall; profiles/r01i_k4.txt.)  Each thread owns a register tile of P pixels x Q output
// channels (conv_layer below); 576 threads because the layers then have 936 / 560 / 480 / 288 tiles per 8 clips.
// Activations are H,W,C with every pixel's channel vector padded by 4 floats, so the 128-bit reads of a warp spread over all
// banks; weights are re-laid [ky][kx][c][q4][g][4] at load so a warp's weight read is one contiguous 128-byte line.
// Why FFMA and not tensor cores: at fp32 accuracy the alternatives are 3xTF32 / 3xBF16 splits on mma.sync; 0.78 MMAC per
// clip puts this kernel at ~1/3 of K1's time per clip already.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "edison_internal.h"
#include "../../include/edison_b200.h"

namespace edison {

constexpr int kCubeClips = 8;        // clips per CTA tile (even: the unpooled layers pair clip b with b + 4)
constexpr int kCubeThreads = 576;         // 18 warps: conv2 has 560 register tiles, conv3 480, conv1 936
constexpr int kPad = 4;              // floats appended to every pixel's channel vector

// activation geometry (H, W, C) after each layer
constexpr int kA1 = 13 * 9 * (16 + kPad);     // conv1 + pool
constexpr int kA2 = 5 * 7 * (32 + kPad);      // conv2 + pool
constexpr int kA3 = 3 * 5 * (64 + kPad);      // conv3
constexpr int kA4 = 1 * 3 * 32;               // conv4 (flattened H,W,C = dense input, unpadded)
constexpr int kBufA = kA2;                    // conv2 out / conv4 out
constexpr int kBufB = kA1;                    // conv1 out / conv3 out
constexpr int kW1 = 5 * 5 * 1 * 16, kW2 = 3 * 3 * 16 * 32, kW3 = 3 * 3 * 32 * 64, kW4 = 3 * 3 * 64 * 32;
static_assert(kW3 == kW4 && 3 * kA4 <= kBufA && kA3 <= kBufB, "buffer sharing");

struct CubeModel {
  float w1[kW1], b1[16];
  float w2[kW2], b2[32];
  float w3[kW3], b3[64];
  float w4[kW4], b4[32];
  float wd[10 * 96], bd[10];         // dense stays [out][in]
  float pad_[2];
};

struct CubeSmem {
  float wbig[kW3];                            // w3 during conv3, w4 during conv4
  float w1[kW1], w2[kW2];
  float stage[kCubeClips * EDISON_NET_IN];    // the tile's input, contiguous as in HBM
  float a[kCubeClips][kBufA];
  float b[kCubeClips][kBufB];
  float logits[kCubeClips][12];
};
static_assert(sizeof(CubeSmem) <= 227 * 1024, "tile + weights exceed the shared memory of an SM");

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
}
// n floats (a multiple of 4, both sides 16-byte aligned) by the whole CTA
__device__ __forceinline__ void stage_floats(float *dst, const float *src, int n) {
  for (int i = threadIdx.x * 4; i < n; i += kCubeThreads * 4) cp_async16(dst + i, src + i);
}

// One conv layer over the tile.  in: [clip][HIN][WIN][C + pad] (C == 1: unpadded scalars), out: [clip][HP][WO][O + opad];
// wgt in shared memory.  A thread's register tile is P pixels x Q output channels, P = PB clips (b, b + 8/PB, ...) times the
// two conv rows under a pooled pixel (POOL) or one pixel; S neighbouring lanes split the input channels of the same tile
// and add their sums by shuffle; KS > 1 deals the kernel rows to KS threads, each writing a raw partial sum (no bias, no
// ReLU) to out[clip][split][pixel][O] for the consumer to add up.  Shared memory hands a lane 4 B per clock whatever the broadcast, so the tile shape sets
// the ceiling: P*Q FFMAs per P+Q loaded floats (2x8 = 1.6 held the FMA pipe at 41 %; 4x8 = 2.7, 8x8 = 4.0).
template <int HIN, int WIN, int C, int KH, int KW, int O, int Q, bool POOL, int PB, int S, int KS, bool PIXBIT, int IN_STRIDE, int OUT_STRIDE, int OPAD>
__device__ __forceinline__ void conv_layer(const float *__restrict__ in, float *__restrict__ out, const float *__restrict__ wgt,
                                           const float *__restrict__ bias) {
  constexpr int HO = HIN - KH + 1, WO = WIN - KW + 1;
  constexpr int HP = POOL ? HO / 2 : HO;
  constexpr int G = O / Q;
  constexpr int CP = C == 1 ? 1 : C + kPad;
  constexpr int OP = O + OPAD;
  constexpr int NPIX = HP * WO;
  constexpr int NB = kCubeClips / PB;            // clip groups; a thread's clips are b, b + NB, ...
  constexpr int R = POOL ? 2 : 1;                // conv rows per clip in the tile
  constexpr int P = PB * R;
  constexpr int NREST = KS * NB * NPIX;                            // (row split, clip group, pixel) triples
  static_assert(!PIXBIT || KS == 1 || (NB * NPIX) % 2 == 0, "a lane pair must not straddle two kernel rows");
  constexpr int NITEMS = S * G * (PIXBIT ? (NREST + 1) / 2 * 2 : NREST);
  constexpr int KROWS = KH / KS;
  constexpr int C4 = C == 1 ? 1 : C / 4;
  constexpr int LB = PIXBIT ? 2 : 1;              // lanes of one (g, s) before the next
  static_assert(kCubeClips % PB == 0 && (S & (S - 1)) == 0 && LB * S * G <= 32 && 32 % (LB * S * G) == 0 && (C == 1 ? S == 1 : C4 % S == 0), "tile shape");
  static_assert(KH % KS == 0 && (KS == 1 || !POOL), "row split");
  constexpr int NROUNDS = (NITEMS + kCubeThreads - 1) / kCubeThreads;
#pragma unroll 1
  for (int round = 0; round < NROUNDS; round++) {
    const int item_raw = round * kCubeThreads + threadIdx.x;
    if (S > 1 ? (item_raw & ~31) >= NITEMS : item_raw >= NITEMS) break;      // whole warps stay for the shuffles
    // lane order: [pixel bit] g s rest.  A 128-bit shared load costs the data pipe 4 clocks when neighbouring lanes read
    // different addresses and 2.1 when each even/odd pair shares one or the pairs alternate between two (tools/ubench/lds.cu):
    // with the pixel bit lowest, weight reads are of the first kind of cheap and pixel reads of the second.
    const int item = item_raw < NITEMS ? item_raw : 0;
    const int g = (item / LB) % G;
    const int s = (item / (LB * G)) % S;
    int rest = PIXBIT ? 2 * (item / (2 * S * G)) + (item & 1) : item / (S * G);
    const bool valid = item_raw < NITEMS && rest < NREST;
    if (rest >= NREST) rest = 0;
    const int ks = KS > 1 ? rest / (NB * NPIX) : 0;       // slowest, so that a lane pair reads the same kernel row's weights
    if (KS > 1) rest -= ks * (NB * NPIX);
    const int b = rest / NPIX, pix = rest - b * NPIX;
    const int py = pix / WO, x = pix - py * WO;
    const float *p0 = in + b * IN_STRIDE + ((R * py) * WIN + x) * CP + (C == 1 ? 0 : 4 * s);
    float acc[P][Q];
#pragma unroll
    for (int p = 0; p < P; p++)
#pragma unroll
      for (int q = 0; q < Q; q++) acc[p][q] = 0.f;
    const float *wg = wgt + (s * G + g) * 4;     // weights [k][c4 / S][cc][q4][s][g][4]: a warp's read is contiguous lines
#pragma unroll 1
    for (int ky = ks * KROWS; ky < (ks + 1) * KROWS; ky++) {
      if (C == 1) {
#pragma unroll
        for (int kx = 0; kx < KW; kx++) {
          float av[P];
#pragma unroll
          for (int p = 0; p < P; p++) av[p] = p0[(p / R) * NB * IN_STRIDE + (p % R) * WIN + ky * WIN + kx];
          const float *wk = wg + (ky * KW + kx) * O;
#pragma unroll
          for (int q4 = 0; q4 < Q / 4; q4++) {
            const float4 w = *reinterpret_cast<const float4 *>(wk + q4 * G * 4);
#pragma unroll
            for (int p = 0; p < P; p++) {
              acc[p][4 * q4 + 0] = fmaf(av[p], w.x, acc[p][4 * q4 + 0]);
              acc[p][4 * q4 + 1] = fmaf(av[p], w.y, acc[p][4 * q4 + 1]);
              acc[p][4 * q4 + 2] = fmaf(av[p], w.z, acc[p][4 * q4 + 2]);
              acc[p][4 * q4 + 3] = fmaf(av[p], w.w, acc[p][4 * q4 + 3]);
            }
          }
        }
      } else {
#pragma unroll 1
        for (int kx = 0; kx < KW; kx++) {
          const float *q0 = p0 + (ky * WIN + kx) * CP;
          const float *wk = wg + (ky * KW + kx) * C * O;
#pragma unroll
          for (int c4 = 0; c4 < C4 / S; c4++) {
            float4 v[P];
#pragma unroll
            for (int p = 0; p < P; p++)
              v[p] = *reinterpret_cast<const float4 *>(q0 + (p / R) * NB * IN_STRIDE + (p % R) * WIN * CP + 4 * S * c4);
#pragma unroll
            for (int cc = 0; cc < 4; cc++) {
              const float *wr = wk + (4 * c4 + cc) * S * O;
#pragma unroll
              for (int q4 = 0; q4 < Q / 4; q4++) {
                const float4 w = *reinterpret_cast<const float4 *>(wr + q4 * S * G * 4);
#pragma unroll
                for (int p = 0; p < P; p++) {
                  const float a = cc == 0 ? v[p].x : cc == 1 ? v[p].y : cc == 2 ? v[p].z : v[p].w;
                  acc[p][4 * q4 + 0] = fmaf(a, w.x, acc[p][4 * q4 + 0]);
                  acc[p][4 * q4 + 1] = fmaf(a, w.y, acc[p][4 * q4 + 1]);
                  acc[p][4 * q4 + 2] = fmaf(a, w.z, acc[p][4 * q4 + 2]);
                  acc[p][4 * q4 + 3] = fmaf(a, w.w, acc[p][4 * q4 + 3]);
                }
              }
            }
          }
        }
      }
    }
    if (S > 1) {
#pragma unroll
      for (int p = 0; p < P; p++)
#pragma unroll
        for (int q = 0; q < Q; q++)
#pragma unroll
          for (int m = 1; m < S; m <<= 1) acc[p][q] += __shfl_xor_sync(0xffffffffu, acc[p][q], m * G * LB);
    }
    // bias, ReLU (nl_func_relu_array_f32), then max pool over the two rows (pool_func_mp_array_f32); lane s of a split
    // group stores clips s, s + S, ... of the tile
    float bq[Q];
#pragma unroll
    for (int q = 0; q < Q; q++) bq[q] = KS > 1 ? 0.f : __ldg(bias + g * Q + q);
#pragma unroll
    for (int j = 0; j < PB; j++) {
      if (!valid || (S > 1 && (j % S) != s)) continue;
      float *o = out + (b + j * NB) * OUT_STRIDE + ((ks * HP + py) * WO + x) * OP + g * Q;
#pragma unroll
      for (int q4 = 0; q4 < Q / 4; q4++) {
        float r[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const float t = POOL ? fmaxf(acc[j * R][4 * q4 + i], acc[j * R + R - 1][4 * q4 + i]) : acc[j][4 * q4 + i];
          r[i] = KS > 1 ? t : fmaxf(t + bq[4 * q4 + i], 0.f);
        }
        *reinterpret_cast<float4 *>(o + 4 * q4) = make_float4(r[0], r[1], r[2], r[3]);
      }
    }
  }
}

// the tile's input: 16-byte copies when the tile is whole and aligned, else word copies + zero fill
__device__ __forceinline__ void stage_input(float *stage, const float *net_in, long long c0, long long n, bool aligned) {
  const int nc = (int)((n - c0) < kCubeClips ? (n - c0) : kCubeClips);
  const float *src = net_in + c0 * EDISON_NET_IN;
  if (nc == kCubeClips && aligned) {
    stage_floats(stage, src, kCubeClips * EDISON_NET_IN);
  } else {
    for (int i = threadIdx.x; i < kCubeClips * EDISON_NET_IN; i += kCubeThreads) {
      if (i < nc * EDISON_NET_IN) cp_async4(stage + i, src + i);
      else stage[i] = 0.f;
    }
  }
}

__global__ void __launch_bounds__(kCubeThreads, 1)
cube_f32_kernel(const CubeModel *__restrict__ M, const float *__restrict__ net_in, long long n, float *__restrict__ logits,
                float *__restrict__ softmax) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  CubeSmem &S = *reinterpret_cast<CubeSmem *>(smem_raw);
  const long long n_tiles = (n + kCubeClips - 1) / kCubeClips;
  const bool aligned = (reinterpret_cast<unsigned long long>(net_in) & 15) == 0;   // a tile is 12,896 B: alignment is the base's
  if ((long long)blockIdx.x >= n_tiles) return;
  stage_floats(S.w1, M->w1, kW1);
  stage_floats(S.w2, M->w2, kW2);
  stage_floats(S.wbig, M->w3, kW3);
  stage_input(S.stage, net_in, (long long)blockIdx.x * kCubeClips, n, aligned);
  cp_async_wait_all();
  __syncthreads();
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long c0 = tile * kCubeClips;
    const int nc = (int)((n - c0) < kCubeClips ? (n - c0) : kCubeClips);
    conv_layer<31, 13, 1, 5, 5, 16, 8, true, 2, 1, 1, false, EDISON_NET_IN, kBufB, kPad>(S.stage, &S.b[0][0], S.w1, M->b1);
    __syncthreads();
    if (tile + gridDim.x < n_tiles) stage_input(S.stage, net_in, (tile + gridDim.x) * kCubeClips, n, aligned);   // lands during conv2..4
    conv_layer<13, 9, 16, 3, 3, 32, 8, true, 2, 1, 1, true, kBufB, kBufA, kPad>(&S.b[0][0], &S.a[0][0], S.w2, M->b2);
    cp_async_wait_all();                  // w3 (issued after the previous tile's conv4) and the next input
    __syncthreads();
    conv_layer<5, 7, 32, 3, 3, 64, 8, false, 4, 2, 1, true, kBufA, kBufB, kPad>(&S.a[0][0], &S.b[0][0], S.wbig, M->b3);
    __syncthreads();
    stage_floats(S.wbig, M->w4, kW4);
    cp_async_wait_all();
    __syncthreads();
    conv_layer<3, 5, 64, 3, 3, 32, 4, false, 4, 2, 3, true, kBufB, kBufA, 0>(&S.b[0][0], &S.a[0][0], S.wbig, M->b4);
    __syncthreads();
    if (tile + gridDim.x < n_tiles) stage_floats(S.wbig, M->w3, kW3);     // lands during the next tile's conv1/conv2
    // conv4 bias + ReLU on the sum of its three partial sums, then dense (forward_dense): 4 lanes per (clip, class)
    if (threadIdx.x < kCubeClips * EDISON_N_CLASSES * 4) {
      const int part = threadIdx.x & 3, bo = threadIdx.x >> 2;
      const int b = bo / EDISON_N_CLASSES, o = bo - b * EDISON_N_CLASSES;
      const float *x = &S.a[b][0], *w = M->wd + o * 96;
      float acc = part == 0 ? __ldg(M->bd + o) : 0.f;
#pragma unroll 8
      for (int j = part * 24; j < part * 24 + 24; j++) {
        const float v = fmaxf(__ldg(M->b4 + (j & 31)) + x[j] + x[kA4 + j] + x[2 * kA4 + j], 0.f);
        acc = fmaf(v, __ldg(w + j), acc);
      }
      acc += __shfl_xor_sync(0xffffffffu, acc, 1);
      acc += __shfl_xor_sync(0xffffffffu, acc, 2);
      if (part == 0) S.logits[b][o] = acc;
    }
    __syncthreads();
    // softmax, natural e (forward_sm)
    if (threadIdx.x < nc) {
      const int b = threadIdx.x;
      float mx = S.logits[b][0];
#pragma unroll
      for (int o = 1; o < EDISON_N_CLASSES; o++) mx = fmaxf(mx, S.logits[b][o]);
      float e[EDISON_N_CLASSES], sum = 0.f;
#pragma unroll
      for (int o = 0; o < EDISON_N_CLASSES; o++) { e[o] = expf(S.logits[b][o] - mx); sum += e[o]; }
      const float inv = 1.0f / sum;
#pragma unroll
      for (int o = 0; o < EDISON_N_CLASSES; o++) {
        if (logits) logits[(c0 + b) * EDISON_N_CLASSES + o] = S.logits[b][o];
        if (softmax) softmax[(c0 + b) * EDISON_N_CLASSES + o] = e[o] * inv;
      }
    }
  }
}

// ---- model load ----------------------------------------------------------------------------------------------------------
static CubeModel *g_cube = nullptr;   // device; loaded on first use under the ctx lock
static bool g_cube_cfg = false;

static int cube_load(const char *path) {
  FILE *f = fopen(path, "rb");
  if (!f) return ctx_fail(EDISON_E_DATA, "cannot open %s", path);
  std::vector<unsigned char> raw;
  unsigned char buf[65536];
  size_t got;
  while ((got = fread(buf, 1, sizeof buf, f)) > 0) raw.insert(raw.end(), buf, buf + got);
  fclose(f);
  if (raw.size() < 16 || memcmp(raw.data(), "EDCUBEF\0", 8) != 0) return ctx_fail(EDISON_E_DATA, "%s: not a float model blob", path);
  uint32_t ver, nl;
  memcpy(&ver, &raw[8], 4); memcpy(&nl, &raw[12], 4);
  if (ver != 1 || nl != 5) return ctx_fail(EDISON_E_MODEL, "%s: version %u / %u layers unsupported", path, ver, nl);
  static const uint32_t want[5][7] = {{1, 16, 5, 5, 1, 1, 2}, {1, 32, 3, 3, 16, 1, 2}, {1, 64, 3, 3, 32, 1, 0},
                                      {1, 32, 3, 3, 64, 1, 0}, {2, 10, 1, 1, 96, 0, 0}};
  CubeModel *h = new CubeModel;
  memset(h, 0, sizeof *h);
  float *wdst[5] = {h->w1, h->w2, h->w3, h->w4, h->wd}, *bdst[5] = {h->b1, h->b2, h->b3, h->b4, h->bd};
  size_t off = 16;
  for (int l = 0; l < 5; l++) {
    uint32_t hd[7];
    if (off + 28 > raw.size()) { delete h; return ctx_fail(EDISON_E_DATA, "%s: truncated", path); }
    memcpy(hd, &raw[off], 28); off += 28;
    if (memcmp(hd, want[l], 28) != 0) { delete h; return ctx_fail(EDISON_E_MODEL, "%s: layer %d is not the kws_conv graph", path, l); }
    const size_t co = hd[1], kh = hd[2], kw = hd[3], ci = hd[4], nw = co * kh * kw * ci;
    if (off + 4 * (nw + co) > raw.size()) { delete h; return ctx_fail(EDISON_E_DATA, "%s: truncated", path); }
    const float *w = reinterpret_cast<const float *>(&raw[off]);
    if (hd[0] == 1) {
      // per-thread channel tile Q and lane split S of each layer (conv_layer<..., Q, ..., S, ...>)
      const size_t Q = l == 3 ? 4 : 8, S = l == 2 ? 2 : l == 3 ? 2 : 1, G = co / Q;
      for (size_t o = 0; o < co; o++) {                // O,H,W,I (kws.c strides) -> [k][c4 / S][cc][q4][s][g][4], o = g*Q + 4*q4 + j
        const size_t g = o / Q, q4 = (o % Q) / 4, j = o % 4;
        for (size_t k = 0; k < kh * kw; k++)
          for (size_t c = 0; c < ci; c++) {
            const size_t c4 = c / 4, cc = c % 4, sp = ci == 1 ? 0 : c4 % S, jj = ci == 1 ? 0 : c4 / S;
            const size_t row = ci == 1 ? k : (k * (ci / 4 / S) + jj) * 4 + cc;        // (tap, channel-block group, cc)
            wdst[l][((row * (Q / 4) + q4) * S + sp) * G * 4 + g * 4 + j] = w[(o * kh * kw + k) * ci + c];
          }
      }
    } else {
      memcpy(wdst[l], w, 4 * nw);
    }
    off += 4 * nw;
    memcpy(bdst[l], &raw[off], 4 * co);
    off += 4 * co;
  }
  CubeModel *d = nullptr;
  cudaError_t e = cudaMalloc(&d, sizeof *h);
  if (e == cudaSuccess) e = cudaMemcpy(d, h, sizeof *h, cudaMemcpyHostToDevice);
  delete h;
  if (e != cudaSuccess) { if (d) cudaFree(d); return ctx_fail(EDISON_E_CUDA, "float model upload: %s", cudaGetErrorString(e)); }
  g_cube = d;
  return EDISON_OK;
}

static int cube_launch(const CtxView &c, const float *d_in, size_t n, float *d_logits, float *d_softmax, cudaStream_t st) {
  if (!g_cube_cfg) {
    cudaError_t e = cudaFuncSetAttribute(cube_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(CubeSmem));
    if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "cube_f32 shared memory opt-in: %s", cudaGetErrorString(e));
    g_cube_cfg = true;
  }
  if (!g_cube) {
    char path[1024];
    snprintf(path, sizeof path, "%s/kws_cube_f32.bin", c.data_dir);
    int rc = cube_load(path);
    if (rc) return rc;
  }
  if (!n) return EDISON_OK;
  const long long tiles = ((long long)n + kCubeClips - 1) / kCubeClips;
  const unsigned grid = (unsigned)(tiles < c.sm_count ? tiles : c.sm_count);
  cube_f32_kernel<<<grid, kCubeThreads, sizeof(CubeSmem), st>>>(g_cube, d_in, (long long)n, d_logits, d_softmax);
  count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "cube_f32_kernel: %s", cudaGetErrorString(e));
  return EDISON_OK;
}

void cube_release(void) {
  if (g_cube) cudaFree(g_cube);
  g_cube = nullptr;
}

static bool cube_dev_ptr(const void *p) {
  cudaPointerAttributes a;
  if (!p || cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

}  // namespace edison

using namespace edison;

extern "C" {

int edison_infer_f32_batch_dev(const float *d_net_in, size_t n, float *d_logits, float *d_softmax, void *stream) {
  DeviceScope device_scope_;
  if (n && !d_net_in) return ctx_fail(EDISON_E_ARG, "edison_infer_f32_batch_dev: null input");
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  if (!rc) rc = cube_launch(c, d_net_in, n, d_logits, d_softmax, stream ? (cudaStream_t)stream : c.stream);
  ctx_unlock();
  return rc;
}

int edison_infer_f32_batch(const float *net_in, size_t n, float *logits, float *softmax) {
  DeviceScope device_scope_;
  if (n && !net_in) return ctx_fail(EDISON_E_ARG, "edison_infer_f32_batch: null input");
  if (cube_dev_ptr(net_in)) {
    int rc = edison_infer_f32_batch_dev(net_in, n, logits, softmax, nullptr);
    return rc ? rc : edison_sync(nullptr);
  }
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  float *d_in = nullptr, *d_lg = nullptr, *d_sm = nullptr;
  if (!rc && n) {
    cudaError_t e = cudaMalloc(&d_in, n * EDISON_NET_IN * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&d_lg, n * EDISON_N_CLASSES * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&d_sm, n * EDISON_N_CLASSES * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_in, net_in, n * EDISON_NET_IN * sizeof(float), cudaMemcpyHostToDevice, c.stream);
    if (e != cudaSuccess) rc = ctx_fail(EDISON_E_NOMEM, "edison_infer_f32_batch staging: %s", cudaGetErrorString(e));
    if (!rc) rc = cube_launch(c, d_in, n, d_lg, d_sm, c.stream);
    if (!rc) {
      if (logits) e = cudaMemcpyAsync(logits, d_lg, n * EDISON_N_CLASSES * sizeof(float), cudaMemcpyDeviceToHost, c.stream);
      if (e == cudaSuccess && softmax)
        e = cudaMemcpyAsync(softmax, d_sm, n * EDISON_N_CLASSES * sizeof(float), cudaMemcpyDeviceToHost, c.stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
      if (e != cudaSuccess) rc = ctx_fail(EDISON_E_CUDA, "edison_infer_f32_batch: %s", cudaGetErrorString(e));
    }
    if (d_in) cudaFree(d_in);
    if (d_lg) cudaFree(d_lg);
    if (d_sm) cudaFree(d_sm);
  }
  ctx_unlock();
  return rc;
}

}  // extern "C"
