// nnom_q7_kernel.cu -- K2: the int8 NNoM keyword-spotting net (firmware/src/ai/nnom/kws_nnom/weights.h:138-161)
// over a batch of 31x13 inputs.
//
// Replaces model_run (firmware/src/ai/nnom/src/core/nnom.c:996-1041) for this graph:
//   Input(31,13,1) -> Conv 16@5x5 -> ReLU -> MaxPool(2,1) -> Conv 32@3x3 -> ReLU -> MaxPool(2,1)
//   -> Conv 64@3x3 -> ReLU -> Conv 32@3x3 -> ReLU -> Dense 10 -> Softmax
// with the arithmetic of arm_convolve_HWC_q7_{basic,fast}_nonsquare / local_convolve_HWC_q7_nonsquare
// (nnom_local.c:425-484), local_maxpool_q7_HWC (nnom_local.c:117-160), arm_fully_connected_q7_opt
// (weights de-interleaved at conversion time) and the three softmax variants:
//   out = SSAT8( ((bias << bias_lshift) + (1 << (out_rshift-1)) + sum in*w) >> out_rshift )
// int32 accumulation is exact and order independent, so any tiling gives the reference's bits.
//
// This revision keeps a tile of clips in shared memory through all layers (activations never touch
// HBM: 403 B in, 10 + 10 B out per clip) and does the contractions with IDP.4A.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "edison_internal.h"
#include "../../include/edison_b200.h"

namespace edison {

// topology this build has kernels for
constexpr int kInH = 31, kInW = 13;
struct LayerDesc { int cout, kh, kw, cin, rshift, lshift, relu, pool; };
constexpr int kC1 = 16, kC2 = 32, kC3 = 64, kC4 = 32, kDenseIn = 96, kDenseOut = 10;

// packed device model: weights in the reference's O,H,W,I order, int8
struct NnomModel {
  int8_t w1[kC1 * 5 * 5 * 1];
  int8_t w2[kC2 * 3 * 3 * kC1];
  int8_t w3[kC3 * 3 * 3 * kC2];
  int8_t w4[kC4 * 3 * 3 * kC3];
  int8_t wd[kDenseOut * kDenseIn];
  int8_t b1[kC1], b2[kC2], b3[kC3], b4[kC4], bd[16];
  int rs[5], ls[5];
  int pad_[2];
};
static_assert(sizeof(NnomModel) % 16 == 0, "NnomModel is copied with 16-byte vectors");

constexpr int kClipsPerCta = 8;
constexpr int kNetThreads = 256;
constexpr int kBuf0 = 1120;   // input (403 -> 404), conv2 pooled 5*7*32, conv4 1*3*32
constexpr int kBuf1 = 1872;   // conv1 pooled 13*9*16, conv3 3*5*64, logits

struct NetSmem {
  NnomModel M;
  int8_t buf0[kClipsPerCta][kBuf0];
  int8_t buf1[kClipsPerCta][kBuf1];
};

__device__ __forceinline__ int ssat8(int v) { return min(max(v, -128), 127); }
__device__ __forceinline__ int usat(int v, int bits) { return min(max(v, 0), (1 << bits) - 1); }

// valid, stride-1 HWC convolution + ReLU + (POOL x 1) max pool, 4 output channels per work item
template <int H, int W, int C, int O, int KH, int KW, int POOL, int IN_STRIDE, int OUT_STRIDE>
__device__ __forceinline__ void conv_layer(const int8_t *in, int8_t *out, const int8_t *w, const int8_t *bias,
                                           int rs, int ls, int nclips) {
  constexpr int Ho = H - KH + 1, Wo = W - KW + 1;
  constexpr int Hp = POOL > 1 ? (Ho - POOL) / POOL + 1 : Ho;
  constexpr int R = POOL > 1 ? POOL : 1;
  constexpr int OG = O / 4;
  const int items = nclips * Hp * Wo * OG;
  for (int idx = threadIdx.x; idx < items; idx += kNetThreads) {
    const int og = idx % OG;
    int t = idx / OG;
    const int x = t % Wo; t /= Wo;
    const int yp = t % Hp;
    const int clip = t / Hp;
    const int8_t *ib = in + clip * IN_STRIDE;
    int acc[R][4];
#pragma unroll
    for (int r = 0; r < R; r++)
#pragma unroll
      for (int o = 0; o < 4; o++) acc[r][o] = ((int)bias[4 * og + o] << ls) + (1 << (rs - 1));
    if constexpr (C % 4 == 0) {
      constexpr int C4 = C / 4;
      const int *iw = reinterpret_cast<const int *>(ib);
      const int *ww = reinterpret_cast<const int *>(w) + (4 * og) * KH * KW * C4;
#pragma unroll 1
      for (int ky = 0; ky < KH; ky++) {
#pragma unroll
        for (int kx = 0; kx < KW; kx++) {
#pragma unroll
          for (int c = 0; c < C4; c++) {
            int wv[4];
#pragma unroll
            for (int o = 0; o < 4; o++) wv[o] = ww[(o * KH * KW + ky * KW + kx) * C4 + c];
#pragma unroll
            for (int r = 0; r < R; r++) {
              const int a = iw[((yp * R + r + ky) * W + (x + kx)) * C4 + c];
#pragma unroll
              for (int o = 0; o < 4; o++) acc[r][o] = __dp4a(a, wv[o], acc[r][o]);
            }
          }
        }
      }
    } else {
      static_assert(C == 1 || C % 4 == 0, "channel count");
      const int8_t *wb = w + (4 * og) * KH * KW;
#pragma unroll 1
      for (int ky = 0; ky < KH; ky++) {
#pragma unroll
        for (int kx = 0; kx < KW; kx++) {
          int wv[4];
#pragma unroll
          for (int o = 0; o < 4; o++) wv[o] = wb[o * KH * KW + ky * KW + kx];
#pragma unroll
          for (int r = 0; r < R; r++) {
            const int a = ib[(yp * R + r + ky) * W + (x + kx)];
#pragma unroll
            for (int o = 0; o < 4; o++) acc[r][o] += a * wv[o];
          }
        }
      }
    }
    uint32_t packed = 0;
#pragma unroll
    for (int o = 0; o < 4; o++) {
      int v = -128;
#pragma unroll
      for (int r = 0; r < R; r++) v = max(v, ssat8(acc[r][o] >> rs));
      v = max(v, 0);  // ReLU tail (nnom_activation.c:118-126); commutes with the max pool
      packed |= (uint32_t)(v & 0xff) << (8 * o);
    }
    *reinterpret_cast<uint32_t *>(out + clip * OUT_STRIDE + (yp * Wo + x) * O + 4 * og) = packed;
  }
}

// arm_softmax_q7 (reference :217-254, SIMD :83-215) and local_softmax_q7 (nnom_local.c:948-1003)
__device__ void softmax_q7(const int8_t *in, int8_t *out, int mode) {
  constexpr int n = kDenseOut;
  if (mode == EDISON_SOFTMAX_NNOM_LOCAL) {
    int base = -257;
    for (int i = 0; i < n; i++) base = max(base, (int)in[i]);
    base -= 8;
    int sum = 0;
    for (int i = 0; i < n; i++)
      if (in[i] > base) sum += 1 << usat(in[i] - base, 5);
    const int ob = 0x100000 / sum;
    for (int i = 0; i < n; i++) out[i] = in[i] > base ? (int8_t)ssat8(ob >> usat(13 + base - in[i], 5)) : (int8_t)0;
    return;
  }
  int base = -128;
  for (int i = 0; i < n; i++) base = max(base, (int)in[i]);
  base -= 8;
  const int nv = mode == EDISON_SOFTMAX_CMSIS_SIMD ? (n & ~3) : 0;  // handled 4 at a time with 8-bit wrap
  int sum = 0;
  for (int i = 0; i < n; i++) {
    int d = in[i] - base;
    if (i < nv) d = (int)(int8_t)(d & 0xff);  // __SSUB8 wraps (:127)
    sum += 1 << usat(d, 3);
  }
  const int ob = (1 << 20) / sum;
  for (int i = 0; i < n; i++) {
    int d;
    if (i < nv) {
      const int t = (int)(int8_t)((13 - in[i]) & 0xff);                 // __SSUB8(pad, in) :177
      d = (int)(int8_t)((t + (int)(int8_t)(base & 0xff)) & 0xff);       // __SADD8(in, baseV) :178
    } else {
      d = 13 + base - in[i];
    }
    out[i] = (int8_t)ssat8(ob >> usat(d, 5));
  }
}

__global__ void __launch_bounds__(kNetThreads) nnom_q7_kernel(const NnomModel *model, const int8_t *net_in, long long n,
                                                               int8_t *logits, int8_t *softmax, int softmax_mode) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  NetSmem &S = *reinterpret_cast<NetSmem *>(smem_raw);
  {
    const int4 *src = reinterpret_cast<const int4 *>(model);
    int4 *dst = reinterpret_cast<int4 *>(&S.M);
    for (int i = threadIdx.x; i < (int)(sizeof(NnomModel) / 16); i += kNetThreads) dst[i] = src[i];
  }
  __syncthreads();
  const NnomModel &M = S.M;
  const long long n_tiles = (n + kClipsPerCta - 1) / kClipsPerCta;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long c0 = tile * kClipsPerCta;
    const int nclips = (int)min((long long)kClipsPerCta, n - c0);
    for (int i = threadIdx.x; i < nclips * EDISON_NET_IN; i += kNetThreads) {
      const int clip = i / EDISON_NET_IN, j = i - clip * EDISON_NET_IN;
      S.buf0[clip][j] = net_in[(c0 + clip) * EDISON_NET_IN + j];
    }
    __syncthreads();
    conv_layer<31, 13, 1, kC1, 5, 5, 2, kBuf0, kBuf1>(&S.buf0[0][0], &S.buf1[0][0], M.w1, M.b1, M.rs[0], M.ls[0], nclips);
    __syncthreads();
    conv_layer<13, 9, kC1, kC2, 3, 3, 2, kBuf1, kBuf0>(&S.buf1[0][0], &S.buf0[0][0], M.w2, M.b2, M.rs[1], M.ls[1], nclips);
    __syncthreads();
    conv_layer<5, 7, kC2, kC3, 3, 3, 0, kBuf0, kBuf1>(&S.buf0[0][0], &S.buf1[0][0], M.w3, M.b3, M.rs[2], M.ls[2], nclips);
    __syncthreads();
    conv_layer<3, 5, kC3, kC4, 3, 3, 0, kBuf1, kBuf0>(&S.buf1[0][0], &S.buf0[0][0], M.w4, M.b4, M.rs[3], M.ls[3], nclips);
    __syncthreads();
    // Dense 96 -> 10 (nnom_dense.c:92-119), logits into buf1
    for (int idx = threadIdx.x; idx < nclips * kDenseOut; idx += kNetThreads) {
      const int clip = idx / kDenseOut, r = idx - clip * kDenseOut;
      const int *a = reinterpret_cast<const int *>(&S.buf0[clip][0]);
      const int *wr = reinterpret_cast<const int *>(M.wd + r * kDenseIn);
      int acc = ((int)M.bd[r] << M.ls[4]) + (1 << (M.rs[4] - 1));
#pragma unroll
      for (int j = 0; j < kDenseIn / 4; j++) acc = __dp4a(a[j], wr[j], acc);
      const int8_t v = (int8_t)ssat8(acc >> M.rs[4]);
      S.buf1[clip][r] = v;
      if (logits) logits[(c0 + clip) * kDenseOut + r] = v;
    }
    __syncthreads();
    if (softmax && threadIdx.x < nclips) {
      int8_t sm[kDenseOut];
      softmax_q7(&S.buf1[threadIdx.x][0], sm, softmax_mode);
      for (int i = 0; i < kDenseOut; i++) softmax[(c0 + threadIdx.x) * kDenseOut + i] = sm[i];
    }
    __syncthreads();
  }
}

cudaError_t nnom_configure(void) {
  return cudaFuncSetAttribute(nnom_q7_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(NetSmem));
}

cudaError_t launch_nnom_q7(const NnomModel *model, const int8_t *net_in, long long n, int8_t *logits, int8_t *softmax,
                           int softmax_mode, int sm_count, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  const long long n_tiles = (n + kClipsPerCta - 1) / kClipsPerCta;
  long long ctas = n_tiles < (long long)sm_count * 3 ? n_tiles : (long long)sm_count * 3;
  nnom_q7_kernel<<<(unsigned)ctas, kNetThreads, sizeof(NetSmem), stream>>>(model, net_in, n, logits, softmax, softmax_mode);
  count_launch();
  return cudaGetLastError();
}

// ---- blob loader (tools/convert_nnom_weights.py) ------------------------------------------------
int nnom_load(const char *path, NnomModel **out_dev, int *in_scale, int *in_min, int *in_max, char *err, size_t errlen) {
  FILE *fh = fopen(path, "rb");
  if (!fh) { snprintf(err, errlen, "cannot open %s", path); return EDISON_E_DATA; }
  fseek(fh, 0, SEEK_END);
  long sz = ftell(fh);
  fseek(fh, 0, SEEK_SET);
  unsigned char *blob = (unsigned char *)malloc((size_t)sz);
  if (!blob) { fclose(fh); return EDISON_E_NOMEM; }
  if (fread(blob, 1, (size_t)sz, fh) != (size_t)sz) { fclose(fh); free(blob); snprintf(err, errlen, "%s: short read", path); return EDISON_E_DATA; }
  fclose(fh);
  if (sz < 16 || memcmp(blob, "EDNNQ7\0\0", 8)) { free(blob); snprintf(err, errlen, "%s: bad header", path); return EDISON_E_DATA; }
  uint32_t ver, nl;
  memcpy(&ver, blob + 8, 4); memcpy(&nl, blob + 12, 4);
  if (ver != 1) { free(blob); snprintf(err, errlen, "%s: version %u", path, ver); return EDISON_E_DATA; }
  // the graph this build has kernels for: kind, cout, kh, kw, cin, relu, pool
  static const uint32_t expect[5][7] = {{1, kC1, 5, 5, 1, 1, 2}, {1, kC2, 3, 3, kC1, 1, 2}, {1, kC3, 3, 3, kC2, 1, 0},
                                        {1, kC4, 3, 3, kC3, 1, 0}, {2, kDenseOut, 1, 1, kDenseIn, 0, 0}};
  if (nl != 5) { free(blob); snprintf(err, errlen, "%s: %u layers, this build runs the 5-layer kws_conv graph", path, nl); return EDISON_E_MODEL; }
  NnomModel *M = (NnomModel *)calloc(1, sizeof(NnomModel));
  int8_t *wdst[5] = {M->w1, M->w2, M->w3, M->w4, M->wd};
  int8_t *bdst[5] = {M->b1, M->b2, M->b3, M->b4, M->bd};
  const unsigned char *p = blob + 16, *end = blob + sz;
  for (int i = 0; i < 5; i++) {
    uint32_t h[9];
    if (p + 36 > end) { free(blob); free(M); snprintf(err, errlen, "%s: truncated", path); return EDISON_E_DATA; }
    memcpy(h, p, 36); p += 36;
    const uint32_t got[7] = {h[0], h[1], h[2], h[3], h[4], h[7], h[8]};
    if (memcmp(got, expect[i], sizeof got)) { free(blob); free(M); snprintf(err, errlen, "%s: layer %d is not the kws_conv layer this build implements", path, i); return EDISON_E_MODEL; }
    if (h[5] < 1 || h[5] > 24 || h[6] > 24) { free(blob); free(M); snprintf(err, errlen, "%s: layer %d shifts out of range", path, i); return EDISON_E_DATA; }
    const size_t nw = (size_t)h[1] * h[2] * h[3] * h[4];
    if (p + nw + h[1] > end) { free(blob); free(M); snprintf(err, errlen, "%s: truncated", path); return EDISON_E_DATA; }
    memcpy(wdst[i], p, nw); p += nw;
    memcpy(bdst[i], p, h[1]); p += h[1];
    p += (size_t)(-(p - blob)) & 3;
    M->rs[i] = (int)h[5]; M->ls[i] = (int)h[6];
  }
  if (p + 24 > end) { free(blob); free(M); snprintf(err, errlen, "%s: truncated trailer", path); return EDISON_E_DATA; }
  int32_t t[3]; uint32_t s[3];
  memcpy(t, p, 12); memcpy(s, p + 12, 12);
  free(blob);
  if (s[0] != (uint32_t)kInH || s[1] != (uint32_t)kInW || s[2] != 1 || t[0] == 0) { free(M); snprintf(err, errlen, "%s: input shape/scale unsupported", path); return EDISON_E_MODEL; }
  *in_scale = t[0]; *in_min = t[1]; *in_max = t[2];
  NnomModel *d = nullptr;
  if (cudaMalloc(&d, sizeof(NnomModel)) != cudaSuccess) { free(M); return EDISON_E_CUDA; }
  if (cudaMemcpy(d, M, sizeof(NnomModel), cudaMemcpyHostToDevice) != cudaSuccess) { cudaFree(d); free(M); return EDISON_E_CUDA; }
  free(M);
  *out_dev = d;
  return EDISON_OK;
}

void nnom_free(NnomModel *m) { if (m) cudaFree(m); }

}  // namespace edison
