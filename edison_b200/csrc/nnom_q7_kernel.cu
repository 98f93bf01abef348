// nnom_q7_kernel.cu -- the model loader of K2 and the INCREMENTAL sliding-window variant of the int8 NNoM keyword-spotting net
// (firmware/src/ai/nnom/kws_nnom/weights.h:138-161).
//
// K2 itself -- the whole 31x13 window through the net -- is nnom_q7_umma.cu (tcgen05.mma.kind::i8, TMEM accumulators).  This file
// keeps (a) nnom_load: blob -> device model + the tcgen05 kernel's weight image, and (b) nnom_q7_incr_kernel, the opt-in
// streaming mode that advances one new row per layer and hop from HBM rings.  (b) still issues
// mma.sync.m16n8k32.s8 (SASS IMMA) on 3-row images; since the tcgen05 kernel the full window is FASTER than it (100,000 streams:
// 0.21 ms against 0.40 ms per hop), so edison_streams_* default to the full window and (b) is kept for its ring design.
//
// Arithmetic of both: arm_convolve_HWC_q7_{basic,fast}_nonsquare / local_convolve_HWC_q7_nonsquare (nnom_local.c:425-484),
// local_maxpool_q7_HWC (nnom_local.c:117-160), arm_fully_connected_q7_opt (weights de-interleaved at conversion time) and the
// three softmax variants:   out = SSAT8( ((bias << bias_lshift) + (1 << (out_rshift-1)) + sum in*w) >> out_rshift )
// int32 accumulation is exact and order independent (|acc| < 2^24), so any MMA tiling gives the reference's bits.
//
// Incremental kernel layout: every layer is D[M x N] = A[M x K] * B[K x N] with M = output pixels of 16 streams, K = (ky, kx, c) in
// the reference's O,H,W,I weight order, N = output channels.  A fragments are 32-bit loads straight from HWC activation buffers
// (implicit im2col; C = 32 / 64 buffers are XOR-swizzled per pixel so those loads are bank-conflict free), B fragments are
// pre-packed in fragment order at model load.  Epilogue, fused: >> out_rshift, SSAT8, ReLU, int8 store in the next layer's layout.
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "edison_internal.h"
#include "../../include/edison_b200.h"
#include "k2_umma_plan.h"
#include "nnom_blob.h"

namespace edison {
using k2::softmax_q7;
using k2::ssat8;
using k2::usat;

constexpr int kInH = 31, kInW = 13;
constexpr int kC1 = 16, kC2 = 32, kC3 = 64, kC4 = 32, kDenseIn = 96, kDenseOut = 10;

// device model: B fragments [k-step][n-tile][lane] = (b0, b1) of mma.m16n8k32 and accumulator seeds
struct NnomModel {
  uint2 f1[2][2][32];    // conv1: K = 5 rows x 8 (kx padded 5 -> 8) = 40 -> 64, N = 16
  uint2 f2[5][4][32];    // conv2: K = 9 taps x 16 = 144 -> 160, N = 32
  uint2 f3[9][8][32];    // conv3: K = 9 x 32 = 288, N = 64
  uint2 f4[18][4][32];   // conv4: K = 9 x 64 = 576, N = 32
  uint2 fd[3][2][32];    // dense: K = 96, N = 10 -> 16
  int seed1[16], seed2[32], seed3[64], seed4[32], seedd[16];   // (bias << bias_lshift) + (1 << (rshift-1))
  int rs[5];
  int pad_[3];
};
static_assert(sizeof(NnomModel) % 16 == 0, "NnomModel is copied with 16-byte vectors");

constexpr int kNC = 16;            // clips per CTA tile
constexpr int kNetThreads = 256;
constexpr int kNetWarps = kNetThreads / 32;
constexpr int kInStride = EDISON_NET_IN;   // 403: a tile's 16 inputs stay contiguous, exactly as they lie in HBM (conv1 fetches unaligned)
constexpr int kPix1 = 13 * 9, kPix2 = 5 * 7, kPix3 = 3 * 5, kPix4 = 3;   // pixels per clip after each conv(+pool)
__device__ __forceinline__ void mma_s8(int (&d)[4], const uint32_t (&a)[4], uint2 b) {
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}

// word index of 32-bit word w of pixel P in an HWC buffer with C channels; C = 32 / 64 are swizzled
template <int C> __device__ __forceinline__ int act_word(int P, int w) {
  if (C == 32) return P * 8 + (w ^ (((P >> 2) & 1) << 2));
  if (C == 64) return P * 16 + (w ^ (((P >> 1) & 3) << 2));
  return P * (C / 4) + w;
}

// clamp(acc >> rs) to [0, 127]: SSAT8 followed by ReLU (nnom_activation.c:118-126)
__device__ __forceinline__ int q7_relu(int acc, int rs) { return min(max(acc >> rs, 0), 127); }

// ---- 3x3 valid convolution (+ ReLU, + optional 2x1 max pool) as an implicit GEMM -----------------------
// in : HWC int8, pixel index P = clip * PIXS_IN + y * W_IN + x, layout act_word<C>
// out: pixel index Q = clip * (HP * WO) + yo * WO + x, layout act_word<O> when OUT_SWZ else plain HWC
// POOL: an m16 tile = 8 pooled pixels; row g is conv row 2 yo, row g + 8 is conv row 2 yo + 1.
template <int W_IN, int PIXS_IN, int C, int O, int HP, int WO, bool POOL, bool OUT_SWZ, int NSPLIT>
__device__ __forceinline__ void conv3x3_mma(const int8_t *in, int8_t *out, const uint2 *frag, const int *seed, int rs,
                                            int nclips) {
  constexpr int NT = O / 8, NTG = NT / NSPLIT;
  constexpr int KSTEPS = (9 * C + 31) / 32;
  constexpr int PIXS_OUT = HP * WO;
  constexpr int PER_TILE = POOL ? 8 : 16;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  const int rows_total = nclips * PIXS_OUT;
  const int n_items = ((rows_total + PER_TILE - 1) / PER_TILE) * NSPLIT;
  const uint32_t *inw = reinterpret_cast<const uint32_t *>(in);
  for (int item = warp; item < n_items; item += kNetWarps) {
    const int tile = item / NSPLIT, nt0 = (item - tile * NSPLIT) * NTG;
    const int q0 = tile * PER_TILE + g, q1 = POOL ? q0 : q0 + 8;
    const bool v0 = q0 < rows_total, v1 = q1 < rows_total;
    const int qa = v0 ? q0 : 0, qb = v1 ? q1 : 0;
    const int ca = qa / PIXS_OUT, ra = qa - ca * PIXS_OUT, ya = ra / WO, xa = ra - ya * WO;
    const int cb = qb / PIXS_OUT, rb = qb - cb * PIXS_OUT, yb = rb / WO, xb = rb - yb * WO;
    const int Pa = ca * PIXS_IN + (POOL ? 2 * ya : ya) * W_IN + xa;
    const int Pb = cb * PIXS_IN + (POOL ? 2 * yb + 1 : yb) * W_IN + xb;
    int acc[NTG][4];
#pragma unroll
    for (int n = 0; n < NTG; n++) {
      acc[n][0] = acc[n][2] = seed[8 * (nt0 + n) + 2 * t];
      acc[n][1] = acc[n][3] = seed[8 * (nt0 + n) + 2 * t + 1];
    }
#pragma unroll
    for (int ks = 0; ks < KSTEPS; ks++) {
      uint32_t a[4];
      if (C == 16) {  // two taps per k-step; the 10th (padding) tap has zero weights
        const int tap0 = 2 * ks, tap1 = 2 * ks + 1 < 9 ? 2 * ks + 1 : 8;
        const int o0 = (tap0 / 3) * W_IN + tap0 % 3, o1 = (tap1 / 3) * W_IN + tap1 % 3;
        a[0] = inw[act_word<C>(Pa + o0, t)]; a[1] = inw[act_word<C>(Pb + o0, t)];
        a[2] = inw[act_word<C>(Pa + o1, t)]; a[3] = inw[act_word<C>(Pb + o1, t)];
      } else {
        const int tap = C == 32 ? ks : ks >> 1, hb = C == 32 ? 0 : (ks & 1) * 8;
        const int o0 = (tap / 3) * W_IN + tap % 3;
        a[0] = inw[act_word<C>(Pa + o0, hb + t)];     a[1] = inw[act_word<C>(Pb + o0, hb + t)];
        a[2] = inw[act_word<C>(Pa + o0, hb + 4 + t)]; a[3] = inw[act_word<C>(Pb + o0, hb + 4 + t)];
      }
#pragma unroll
      for (int n = 0; n < NTG; n++) mma_s8(acc[n], a, frag[(ks * NT + nt0 + n) * 32 + lane]);
    }
    // epilogue: channels 8 nt + 2t, 2t + 1 of this thread's one (POOL) or two pixels
#pragma unroll
    for (int n = 0; n < NTG; n++) {
      const int w = 2 * (nt0 + n) + (t >> 1), hoff = t & 1;
      if (POOL) {
        const int lo = q7_relu(max(acc[n][0], acc[n][2]), rs), hi = q7_relu(max(acc[n][1], acc[n][3]), rs);
        if (v0) {
          const int wi = OUT_SWZ ? act_word<O>(q0, w) : q0 * (O / 4) + w;
          reinterpret_cast<uint16_t *>(out)[2 * wi + hoff] = (uint16_t)(lo | (hi << 8));
        }
      } else {
        if (v0) {
          const int wi = OUT_SWZ ? act_word<O>(q0, w) : q0 * (O / 4) + w;
          reinterpret_cast<uint16_t *>(out)[2 * wi + hoff] = (uint16_t)(q7_relu(acc[n][0], rs) | (q7_relu(acc[n][1], rs) << 8));
        }
        if (v1) {
          const int wi = OUT_SWZ ? act_word<O>(q1, w) : q1 * (O / 4) + w;
          reinterpret_cast<uint16_t *>(out)[2 * wi + hoff] = (uint16_t)(q7_relu(acc[n][2], rs) | (q7_relu(acc[n][3], rs) << 8));
        }
      }
    }
  }
}

// ---- conv1: 5x5, one input channel, 16 outputs, ReLU, 2x1 max pool ---------------------------------------
// K is re-indexed as ky * 8 + kx (kx padded 5 -> 8, ky padded 5 -> 8 with zero weights), so that the four
// k of one A register are four consecutive input bytes of one row (an unaligned 32-bit fetch).
__device__ __forceinline__ uint32_t ld_unaligned32(const uint32_t *words, int byte_addr) {
  const int wa = byte_addr >> 2;
  return __funnelshift_r(words[wa], words[wa + 1], (byte_addr & 3) * 8);
}

// ---- incremental sliding-window inference (app.c:706-719 advances the window one row per hop) ----------
// A window that starts at global row s is  conv1 rows s..s+26 -> pool pairs (s+2i, s+2i+1) -> conv2 rows on pooled rows
// (a, a+2, a+4), a = s+2i -> pool pairs (a, a+2), a = s+4j -> conv3 on (a, a+4, a+8) -> conv4 on c3[s], c3[s+4], c3[s+8].
// All of these are functions of GLOBAL row indices, so one hop (newest input row r, window start s = r-30) adds exactly one new
// row per layer:  c1[r-4],  p1[r-5] = max(c1[r-5], c1[r-4]),  c2[r-9] = f(p1[r-9], p1[r-7], p1[r-5]),
// p2[r-11] = max(c2[r-11], c2[r-9]),  c3[r-19] = f(p2[r-19], p2[r-15], p2[r-11]),  and the output needs
// c3[r-30], c3[r-26], c3[r-22] from earlier hops.  Rows live in HBM rings [slot][stream][row bytes] (slot = index mod depth;
// depths 1, 5, 3, 9, 12 -> 7,392 B per stream); a hop moves ~3 KB per stream and does 184 k MAC instead of 785 k.
// Integer convolutions are position-local and every pooling pair is anchored at the window start, so the result is
// bit-identical to running K2 on the assembled window (tests/test_gpu_streams.py runs both).
constexpr int kIn5Stride = 80;                     // 5 rows x 13 bytes per stream, padded
constexpr int kRowC1 = 9 * kC1, kRowC2 = 7 * kC2, kRowC3 = 5 * kC3;      // 144, 224, 320 bytes
constexpr int kDepthP1 = 5, kDepthC2 = 3, kDepthP2 = 9, kDepthC3 = 12;

struct IncrSmem {
  NnomModel M;
  alignas(16) int8_t in5[kNC * kIn5Stride + 128];   // conv1 reads zero-weight taps past the 5th row
  alignas(16) int8_t c1n[kNC * kRowC1];             // new conv1 row, plain [stream][9][16]
  alignas(16) int8_t c1p[kNC * kRowC1];             // c1[r-5], stored by the previous hop
  alignas(16) int8_t c2o[kNC * kRowC2];             // c2[r-11], stored two hops ago
  alignas(16) int8_t img2[kNC * 3 * kRowC1];        // p1[r-9], p1[r-7], p1[r-5]: [stream][3][9][16]
  alignas(16) int8_t c2n[kNC * kRowC2];             // new conv2 row, plain [stream][7][32]
  alignas(16) int8_t img3[kNC * 3 * kRowC2];        // p2 rows, act_word<32> layout over P = stream*21 + row*7 + x
  alignas(16) int8_t c3n[kNC * kRowC3];             // new conv3 row, plain [stream][5][64]
  alignas(16) int8_t img4[kNC * 3 * kRowC3];        // c3 rows, act_word<64> layout over P = stream*15 + row*5 + x
  alignas(16) int8_t act4[kNC * kDenseIn];
  alignas(16) int8_t lg[kNC * 16];
};

// conv1 on a 5-row image per stream: one output row (9 pixels x 16 channels, ReLU), no pooling; rows g / g+8 of a tile are
// two different pixels.  Same K indexing as conv1_mma (ky * 8 + kx, zero weights beyond 5x5).
__device__ __forceinline__ void conv1_row_mma(const int8_t *in, int8_t *out, const uint2 *frag, const int *seed, int rs, int ns) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  const int rows_total = ns * 9, n_tiles = (rows_total + 15) / 16;
  const uint32_t *inw = reinterpret_cast<const uint32_t *>(in);
  for (int tile = warp; tile < n_tiles; tile += kNetWarps) {
    const int q0 = tile * 16 + g, q1 = q0 + 8;
    const bool v0 = q0 < rows_total, v1 = q1 < rows_total;
    const int qa = v0 ? q0 : 0, qb = v1 ? q1 : 0;
    const int base_a = (qa / 9) * kIn5Stride + qa % 9, base_b = (qb / 9) * kIn5Stride + qb % 9;
    int acc[2][4];
#pragma unroll
    for (int n = 0; n < 2; n++) {
      acc[n][0] = acc[n][2] = seed[8 * n + 2 * t];
      acc[n][1] = acc[n][3] = seed[8 * n + 2 * t + 1];
    }
#pragma unroll
    for (int ks = 0; ks < 2; ks++) {
      uint32_t a[4];
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const int ky = 4 * ks + 2 * h + (t >> 1), kx0 = 4 * (t & 1);
        if (4 * ks + 2 * h >= 5) {   // zero-weight kernel rows
          a[2 * h] = a[2 * h + 1] = 0;
          continue;
        }
        a[2 * h] = ld_unaligned32(inw, base_a + ky * kInW + kx0);
        a[2 * h + 1] = ld_unaligned32(inw, base_b + ky * kInW + kx0);
      }
      mma_s8(acc[0], a, frag[(ks * 2 + 0) * 32 + lane]);
      mma_s8(acc[1], a, frag[(ks * 2 + 1) * 32 + lane]);
    }
#pragma unroll
    for (int n = 0; n < 2; n++) {
      if (v0) reinterpret_cast<uint16_t *>(out)[(q0 * kC1 + 8 * n + 2 * t) >> 1] = (uint16_t)(q7_relu(acc[n][0], rs) | (q7_relu(acc[n][1], rs) << 8));
      if (v1) reinterpret_cast<uint16_t *>(out)[(q1 * kC1 + 8 * n + 2 * t) >> 1] = (uint16_t)(q7_relu(acc[n][2], rs) | (q7_relu(acc[n][3], rs) << 8));
    }
  }
}

__device__ __forceinline__ int ring_slot(long long index, int depth) { return (int)(((index % depth) + depth) % depth); }
__device__ __forceinline__ uint32_t max_s8x4(uint32_t a, uint32_t b) { return __vmaxs4(a, b); }

// window: [n][403] after this hop's push (rows r-30..r).  r = index of the newest row (hop count - 1).
__global__ void __launch_bounds__(kNetThreads, 2)
nnom_q7_incr_kernel(const NnomModel *model, const int8_t *window, int row0, NnomIncrState st, long long n, long long r, int8_t *logits,
                    int8_t *softmax, int softmax_mode) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  IncrSmem &S = *reinterpret_cast<IncrSmem *>(smem_raw);
  {
    const int4 *src = reinterpret_cast<const int4 *>(model);
    int4 *dst = reinterpret_cast<int4 *>(&S.M);
    for (int i = threadIdx.x; i < (int)(sizeof(NnomModel) / 16); i += kNetThreads) dst[i] = src[i];
    for (int i = threadIdx.x; i < (int)sizeof(S.in5) / 4; i += kNetThreads) reinterpret_cast<uint32_t *>(S.in5)[i] = 0;
  }
  const NnomModel &M = S.M;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  const int tid = threadIdx.x;
  // ring slots of this hop
  const int sp1_new = ring_slot(r - 5, kDepthP1), sp1_a = ring_slot(r - 9, kDepthP1), sp1_b = ring_slot(r - 7, kDepthP1);
  const int sc2_new = ring_slot(r - 9, kDepthC2), sc2_old = ring_slot(r - 11, kDepthC2);
  const int sp2_new = ring_slot(r - 11, kDepthP2), sp2_a = ring_slot(r - 19, kDepthP2), sp2_b = ring_slot(r - 15, kDepthP2);
  const int sc3_new = ring_slot(r - 19, kDepthC3);
  const int sc3_0 = ring_slot(r - 30, kDepthC3), sc3_1 = ring_slot(r - 26, kDepthC3), sc3_2 = ring_slot(r - 22, kDepthC3);
  const long long n_tiles = (n + kNC - 1) / kNC;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long c0 = tile * kNC;
    const int ns = (int)min((long long)kNC, n - c0);
    __syncthreads();
    // (1) everything this hop reads from HBM, issued up front as independent 128-bit loads (the kernel is latency bound):
    //     per stream 60 + 28 + 18 + 9 + 14 = 129 chunks: img4 rows 0..2, img3 rows 0..1, img2 rows 0..1, c1[r-5], c2[r-11]
    {
      constexpr int kJobs = 129, kIter = (kNC * kJobs + kNetThreads - 1) / kNetThreads;
      int4 v[kIter];
      int dst[kIter];        // int4 index into the shared struct, -1 = nothing
#pragma unroll
      for (int it = 0; it < kIter; it++) {
        const int j = tid + it * kNetThreads;
        const int s = j / kJobs, k = j - s * kJobs;
        dst[it] = -1;
        if (s < ns) {
          const long long cs = c0 + s;
          const int4 *src;
          if (k < 60) {
            const int row = k / 20, c = k - row * 20, P = s * 15 + row * 5 + (c >> 2);
            const int slot = row == 0 ? sc3_0 : (row == 1 ? sc3_1 : sc3_2);
            src = reinterpret_cast<const int4 *>(st.c3) + ((long long)slot * n + cs) * 20 + c;
            dst[it] = (int)(offsetof(IncrSmem, img4) / 16) + P * 4 + ((c & 3) ^ ((P >> 1) & 3));
          } else if (k < 88) {
            const int kk = k - 60, row = kk / 14, c = kk - row * 14, P = s * 21 + row * 7 + (c >> 1);
            src = reinterpret_cast<const int4 *>(st.p2) + ((long long)(row ? sp2_b : sp2_a) * n + cs) * 14 + c;
            dst[it] = (int)(offsetof(IncrSmem, img3) / 16) + P * 2 + ((c & 1) ^ ((P >> 2) & 1));
          } else if (k < 106) {
            const int kk = k - 88, row = kk / 9, c = kk - row * 9;
            src = reinterpret_cast<const int4 *>(st.p1) + ((long long)(row ? sp1_b : sp1_a) * n + cs) * 9 + c;
            dst[it] = (int)(offsetof(IncrSmem, img2) / 16) + (s * 27 + row * 9) + c;
          } else if (k < 115) {
            src = reinterpret_cast<const int4 *>(st.c1) + cs * 9 + (k - 106);
            dst[it] = (int)(offsetof(IncrSmem, c1p) / 16) + s * 9 + (k - 106);
          } else {
            src = reinterpret_cast<const int4 *>(st.c2) + ((long long)sc2_old * n + cs) * 14 + (k - 115);
            dst[it] = (int)(offsetof(IncrSmem, c2o) / 16) + s * 14 + (k - 115);
          }
          v[it] = __ldg(src);
        }
      }
      // newest five input rows of each stream: window rows 26..30 (65 bytes at an odd offset: byte loads)
      constexpr int kInIter = (kNC * 65 + kNetThreads - 1) / kNetThreads;
      int8_t b[kInIter];
#pragma unroll
      for (int it = 0; it < kInIter; it++) {
        const int i = tid + it * kNetThreads, s = i / 65, j = i - s * 65;
        b[it] = s < ns ? window[(c0 + s) * EDISON_NET_IN + k2::raw_offset(0, 26 + j / 13, row0) + j % 13] : (int8_t)0;   // ring of rows
      }
#pragma unroll
      for (int it = 0; it < kInIter; it++) {
        const int i = tid + it * kNetThreads, s = i / 65, j = i - s * 65;
        if (s < kNC) S.in5[s * kIn5Stride + j] = b[it];
      }
#pragma unroll
      for (int it = 0; it < kIter; it++)
        if (dst[it] >= 0) reinterpret_cast<int4 *>(&S)[dst[it]] = v[it];
    }
    __syncthreads();
    // (2) conv1 row r-4
    conv1_row_mma(S.in5, S.c1n, &M.f1[0][0][0], M.seed1, M.rs[0], ns);
    __syncthreads();
    // (3) p1[r-5] = max(c1[r-5], c1[r-4]) -> img2 row 2 and the ring; c1[r-4] becomes the stored previous row
    for (int i = tid; i < ns * 9; i += kNetThreads) {
      const int s = i / 9, c = i - s * 9;
      const int4 cur = reinterpret_cast<const int4 *>(S.c1n)[i], prv = reinterpret_cast<const int4 *>(S.c1p)[i];
      const int4 m = make_int4((int)max_s8x4(prv.x, cur.x), (int)max_s8x4(prv.y, cur.y), (int)max_s8x4(prv.z, cur.z), (int)max_s8x4(prv.w, cur.w));
      reinterpret_cast<int4 *>(st.c1)[(c0 + s) * 9 + c] = cur;
      reinterpret_cast<int4 *>(st.p1)[((long long)sp1_new * n + c0 + s) * 9 + c] = m;
      reinterpret_cast<int4 *>(S.img2)[s * 27 + 18 + c] = m;
    }
    __syncthreads();
    // (4) conv2 row r-9 on (p1[r-9], p1[r-7], p1[r-5])
    conv3x3_mma<9, 27, kC1, kC2, 1, 7, false, false, 1>(S.img2, S.c2n, &M.f2[0][0][0], M.seed2, M.rs[1], ns);
    __syncthreads();
    // (5) p2[r-11] = max(c2[r-11], c2[r-9]); store c2[r-9]
    for (int i = tid; i < ns * 14; i += kNetThreads) {
      const int s = i / 14, c = i - s * 14, P = s * 21 + 14 + (c >> 1);
      const int4 cur = reinterpret_cast<const int4 *>(S.c2n)[i], old = reinterpret_cast<const int4 *>(S.c2o)[i];
      const int4 m = make_int4((int)max_s8x4(old.x, cur.x), (int)max_s8x4(old.y, cur.y), (int)max_s8x4(old.z, cur.z), (int)max_s8x4(old.w, cur.w));
      reinterpret_cast<int4 *>(st.c2)[((long long)sc2_new * n + c0 + s) * 14 + c] = cur;
      reinterpret_cast<int4 *>(st.p2)[((long long)sp2_new * n + c0 + s) * 14 + c] = m;
      reinterpret_cast<int4 *>(S.img3)[P * 2 + ((c & 1) ^ ((P >> 2) & 1))] = m;
    }
    __syncthreads();
    // (6) conv3 row r-19 on (p2[r-19], p2[r-15], p2[r-11]); conv4 on the three stored c3 rows (independent of (6))
    conv3x3_mma<7, 21, kC2, kC3, 1, 5, false, false, 2>(S.img3, S.c3n, &M.f3[0][0][0], M.seed3, M.rs[2], ns);
    conv3x3_mma<5, 15, kC3, kC4, 1, 3, false, false, 4>(S.img4, S.act4, &M.f4[0][0][0], M.seed4, M.rs[3], ns);
    __syncthreads();
    for (int i = tid; i < ns * 20; i += kNetThreads) {
      const int s = i / 20, c = i - s * 20;
      reinterpret_cast<int4 *>(st.c3)[((long long)sc3_new * n + c0 + s) * 20 + c] = reinterpret_cast<const int4 *>(S.c3n)[i];
    }
    // (7) dense + softmax, as in nnom_q7_kernel
    if (warp < 2) {
      const uint32_t *aw = reinterpret_cast<const uint32_t *>(S.act4);
      const int r0 = g < ns ? g : 0, r1 = g + 8 < ns ? g + 8 : 0;
      int acc[4];
      acc[0] = acc[2] = M.seedd[8 * warp + 2 * t];
      acc[1] = acc[3] = M.seedd[8 * warp + 2 * t + 1];
#pragma unroll
      for (int ks = 0; ks < 3; ks++) {
        uint32_t a[4];
        a[0] = aw[r0 * 24 + 8 * ks + t];     a[1] = aw[r1 * 24 + 8 * ks + t];
        a[2] = aw[r0 * 24 + 8 * ks + 4 + t]; a[3] = aw[r1 * 24 + 8 * ks + 4 + t];
        mma_s8(acc, a, M.fd[ks][warp][lane]);
      }
      const int rs = M.rs[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int clip = g + 8 * (j >> 1), col = 8 * warp + 2 * t + (j & 1);
        if (clip < ns && col < kDenseOut) {
          const int8_t v = (int8_t)ssat8(acc[j] >> rs);
          S.lg[clip * 16 + col] = v;
          if (logits) logits[(c0 + clip) * kDenseOut + col] = v;
        }
      }
    }
    __syncthreads();
    if (softmax && tid < ns) {
      int8_t sm[kDenseOut];
      softmax_q7(&S.lg[tid * 16], sm, softmax_mode);
      for (int i = 0; i < kDenseOut; i++) softmax[(c0 + tid) * kDenseOut + i] = sm[i];
    }
  }
}

size_t nnom_incr_state_bytes(long long n) {
  return (size_t)n * (kRowC1 * (1 + kDepthP1) + kRowC2 * (kDepthC2 + kDepthP2) + kRowC3 * kDepthC3);
}

void nnom_incr_state_bind(NnomIncrState *st, int8_t *base, long long n) {
  st->c1 = base;
  st->p1 = st->c1 + (size_t)n * kRowC1;
  st->c2 = st->p1 + (size_t)n * kRowC1 * kDepthP1;
  st->p2 = st->c2 + (size_t)n * kRowC2 * kDepthC2;
  st->c3 = st->p2 + (size_t)n * kRowC2 * kDepthP2;
}

cudaError_t launch_nnom_q7_incremental(const NnomModel *model, const int8_t *window, int row0, const NnomIncrState &st, long long n,
                                       long long newest_row, int8_t *logits, int8_t *softmax, int softmax_mode, int sm_count,
                                       cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  const long long n_tiles = (n + kNC - 1) / kNC;
  long long ctas = n_tiles < (long long)sm_count * 2 ? n_tiles : (long long)sm_count * 2;
  nnom_q7_incr_kernel<<<(unsigned)ctas, kNetThreads, sizeof(IncrSmem), stream>>>(model, window, row0, st, n, newest_row, logits, softmax,
                                                                                softmax_mode);
  count_launch();
  return cudaGetLastError();
}

cudaError_t nnom_configure(void) {
  cudaError_t e = cudaFuncSetAttribute(nnom_q7_incr_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(IncrSmem));
  if (e != cudaSuccess) return e;
  return nnom_umma_configure();
}

// weight image of the tcgen05 kernel (nnom_q7_umma.cu); one model per process, like the firmware
static uint8_t *g_umma_image = nullptr;

cudaError_t launch_nnom_q7(const NnomModel *model, const K2Launch &kl, int sm_count, cudaStream_t stream) {
  (void)model;
  if (!g_umma_image) return cudaErrorNotReady;
  return launch_nnom_q7_umma(g_umma_image, kl, sm_count, stream);
}

// ---- blob loader (tools/convert_nnom_weights.py) -> device model ----------------------------------------------------
namespace {
// B[k][n] of one layer -> mma.m16n8k32 B fragments: lane (g = lane >> 2, t = lane & 3) holds
// b0 = B[32 ks + 4t .. +3][8 nt + g], b1 = B[32 ks + 16 + 4t .. +3][8 nt + g]
template <class F> void pack_frags(uint2 *dst, int ksteps, int ntiles, F weight) {
  for (int ks = 0; ks < ksteps; ks++)
    for (int nt = 0; nt < ntiles; nt++)
      for (int lane = 0; lane < 32; lane++) {
        const int g = lane >> 2, t = lane & 3;
        uint32_t b[2] = {0, 0};
        for (int half = 0; half < 2; half++)
          for (int i = 0; i < 4; i++)
            b[half] |= (uint32_t)(uint8_t)weight(32 * ks + 16 * half + 4 * t + i, 8 * nt + g) << (8 * i);
        dst[(ks * ntiles + nt) * 32 + lane] = make_uint2(b[0], b[1]);
      }
}
}  // namespace

int nnom_load(const char *path, NnomModel **out_dev, int *in_scale, int *in_min, int *in_max, char *err, size_t errlen) {
  KwsBlob blob;
  int rc = read_kws_blob(path, blob, err, errlen);
  if (rc) return rc;
  const KwsLayer *L = blob.L;
  *in_scale = blob.in_scale; *in_min = blob.in_min; *in_max = blob.in_max;
  NnomModel *M = (NnomModel *)calloc(1, sizeof(NnomModel));
  if (!M) return EDISON_E_NOMEM;
  // fragments of the incremental (sliding-window) kernel
  // conv1: k = ky * 8 + kx (zero outside the 5x5 window); weights [o][ky][kx]
  pack_frags(&M->f1[0][0][0], 2, 2, [&](int k, int n) -> int8_t {
    const int ky = k >> 3, kx = k & 7;
    return (ky < 5 && kx < 5) ? L[0].w[(n * 5 + ky) * 5 + kx] : (int8_t)0;
  });
  // conv2..4: k = (ky * 3 + kx) * C + c == the reference's O,H,W,I order; zero beyond 9 C
  pack_frags(&M->f2[0][0][0], 5, 4, [&](int k, int n) -> int8_t { return k < 144 ? L[1].w[n * 144 + k] : (int8_t)0; });
  pack_frags(&M->f3[0][0][0], 9, 8, [&](int k, int n) -> int8_t { return L[2].w[n * 288 + k]; });
  pack_frags(&M->f4[0][0][0], 18, 4, [&](int k, int n) -> int8_t { return L[3].w[n * 576 + k]; });
  pack_frags(&M->fd[0][0][0], 3, 2, [&](int k, int n) -> int8_t { return n < kDenseOut ? L[4].w[n * kDenseIn + k] : (int8_t)0; });
  int *seeds[5] = {M->seed1, M->seed2, M->seed3, M->seed4, M->seedd};
  const int8_t *w[5], *b[5];
  int rs[5], ls[5];
  for (int i = 0; i < 5; i++) {
    M->rs[i] = (int)L[i].h[5];
    for (uint32_t o = 0; o < L[i].h[1]; o++)
      seeds[i][o] = (int)((uint32_t)(int32_t)L[i].b[o] << L[i].h[6]) + (1 << (L[i].h[5] - 1));
    w[i] = L[i].w; b[i] = L[i].b; rs[i] = (int)L[i].h[5]; ls[i] = (int)L[i].h[6];
  }
  NnomModel *d = nullptr;
  uint8_t *image = nullptr;
  do {
    if (cudaMalloc(&d, sizeof(NnomModel)) != cudaSuccess) { snprintf(err, errlen, "cudaMalloc(model)"); rc = EDISON_E_CUDA; break; }
    if (cudaMemcpy(d, M, sizeof(NnomModel), cudaMemcpyHostToDevice) != cudaSuccess) { snprintf(err, errlen, "cudaMemcpy(model)"); rc = EDISON_E_CUDA; break; }
    rc = nnom_umma_upload(w, b, rs, ls, &image);
    if (rc) { snprintf(err, errlen, "upload of the K2 weight image failed"); break; }
  } while (0);
  free(M);
  if (rc) {
    cudaGetLastError();
    if (d) cudaFree(d);
    return rc;
  }
  if (g_umma_image) cudaFree(g_umma_image);
  g_umma_image = image;
  *out_dev = d;
  return EDISON_OK;
}

void nnom_free(NnomModel *m) {
  if (m) cudaFree(m);
  if (g_umma_image) cudaFree(g_umma_image);
  g_umma_image = nullptr;
}

}  // namespace edison
