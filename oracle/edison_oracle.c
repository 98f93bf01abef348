/* edison_oracle.c -- CPU restatement of the reference's KWS hot path (plain C99).
 *
 * TEST INFRASTRUCTURE ONLY (see edison_oracle.h).  Each function cites the reference
 * file:line (relative to /root/reference) whose arithmetic it restates.  The canonical
 * arithmetic is the firmware's own build: CMSIS-DSP with ARM_MATH_DSP (Cortex-M4 SIMD
 * branches, firmware/Makefile:30), i.e. oracle/_ref/libref_mfcc_simd.so.
 *
 * Parity status: PINNED against oracle/_ref (tests/test_oracle_cpu.py) and the
 * golden vectors in tests/golden/ generated from it.
 *
 * Written from scratch in an unpacked (re, im as separate int32) form; the M4's packed
 * 16-bit instructions are restated by the helpers directly below.
 */
#include "edison_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------
 * Cortex-M4 DSP semantics on one 16-bit lane (ARMv7-M ARM A7.7; emulations arm_math.h:1368-1697)
 * ---------------------------------------------------------------------------------------- */
static inline int32_t sat16(int32_t v) { return v > 32767 ? 32767 : (v < -32768 ? -32768 : v); }
static inline int32_t qadd(int32_t a, int32_t b) { return sat16(a + b); }   /* QADD16 lane  */
static inline int32_t qsub(int32_t a, int32_t b) { return sat16(a - b); }   /* QSUB16 lane  */
static inline int32_t hadd(int32_t a, int32_t b) { return (a + b) >> 1; }   /* SHADD16 lane */
static inline int32_t hsub(int32_t a, int32_t b) { return (a - b) >> 1; }   /* SHSUB16 lane */
static inline int32_t s16(int32_t v) { return (int16_t)(uint16_t)(uint32_t)v; }

typedef struct { int32_t re, im; } cq15;

/* (co + j*(-si)) applied the way the SIMD branch does it:
 *   re' = SMUAD(C, V) >> 16  = (co*Vre + si*Vim) >> 16
 *   im' = SMUSDX(C, V) >> 16 = (co*Vim - si*Vre) >> 16           arm_cfft_radix4_q15.c:252-256
 * 32-bit wrap-around products like the M4 (no saturation in SMUAD/SMUSDX). */
static inline cq15 twiddle(int32_t co, int32_t si, cq15 v)
{
  cq15 r;
  uint32_t a = (uint32_t)(co * v.re) + (uint32_t)(si * v.im);
  uint32_t b = (uint32_t)(co * v.im) - (uint32_t)(si * v.re);
  r.re = s16((int32_t)a >> 16);
  r.im = s16((int32_t)b >> 16);
  return r;
}

/* ------------------------------------------------------------------------------------------
 * Tables.  arm_common_tables.c is missing from the reference tree (.MISSING_LARGE_BLOBS:4);
 * formulas below reproduce the data inside libarm_cortexM4lf_math.a bit for bit
 * (checked in tests/test_oracle_cpu.py against oracle/_ref/tables_manifest.json).
 * ---------------------------------------------------------------------------------------- */
static int16_t tw1024[1536];
static int16_t tw16[24];
static int16_t realA[32], realB[32];
static int16_t mel_w[1024];
static int16_t mel_start[EO_N_MEL], mel_cnt[EO_N_MEL];
static int     mel_n;
static int     inited;

static int16_t q15_floor(double v)
{
  double f = floor(v);
  if (f > 32767.0) f = 32767.0;
  if (f < -32768.0) f = -32768.0;
  return (int16_t)f;
}

static void gen_twiddle(int16_t *t, int n)
{
  /* twiddleCoef_N_q15: cos, sin interleaved, 3N/4 pairs (arm_common_tables.h:386) */
  for (int i = 0; i < 3 * n / 4; i++) {
    t[2 * i]     = q15_floor(cos(2.0 * M_PI * i / n) * 32768.0);
    t[2 * i + 1] = q15_floor(sin(2.0 * M_PI * i / n) * 32768.0);
  }
}

/* mel filter bank: audio/edison/mfcc/mfcc_utils.py:30-73 (gen_mel_weight_matrix) with the
 * constants of audio/config.py:12-36, quantised as mfcc_on_mcu.py:84 (int16(128*W), truncation)
 * and compacted as mfcc_on_mcu.py:28-62 (melMtxToUnspares). */
static double hz_to_mel(double hz) { return 1127.0 * log(1.0 + hz / 700.0); }

static void gen_mel(void)
{
  const int nbins = 513, nmel = EO_N_MEL;
  const double lo = hz_to_mel(80.0), hi = hz_to_mel(7600.0);
  double edges[EO_N_MEL + 2];
  const double step = (hi - lo) / (nmel + 1);
  for (int i = 0; i < nmel + 2; i++) edges[i] = i * step + lo;     /* np.linspace */
  edges[nmel + 1] = hi;
  static int16_t dense[513][EO_N_MEL];
  memset(dense, 0, sizeof dense);
  for (int b = 1; b < nbins; b++) {                                 /* bin 0 (DC) zeroed */
    double f = b * (8000.0 / 512.0);
    if (b == nbins - 1) f = 8000.0;
    double m = hz_to_mel(f);
    for (int k = 0; k < nmel; k++) {
      double ls = (m - edges[k]) / (edges[k + 1] - edges[k]);
      double us = (edges[k + 2] - m) / (edges[k + 2] - edges[k + 1]);
      double w = ls < us ? ls : us;
      if (w < 0.0) w = 0.0;
      dense[b][k] = (int16_t)(128.0 * w);
    }
  }
  mel_n = 0;
  for (int k = 0; k < nmel; k++) {
    int first = -1, cnt = 0;
    for (int b = 0; b < nbins; b++)
      if (dense[b][k] != 0) { if (first < 0) first = b; cnt++; }
    mel_start[k] = (int16_t)first;
    mel_cnt[k] = (int16_t)cnt;
    for (int b = first; b < first + cnt; b++) mel_w[mel_n++] = dense[b][k];
  }
}

int eo_init(void)
{
  if (inited) return 0;
  gen_twiddle(tw1024, 1024);
  gen_twiddle(tw16, 16);
  /* realCoefAQ15 / realCoefBQ15 (arm_common_tables.h:411-412), read at stride 2*modifier = 512
   * q15 by arm_split_rfft_q15 for fftLenReal = 32 (arm_rfft_init_q15.c:213-222) */
  for (int n = 0; n < 16; n++) {
    double th = 2.0 * M_PI * (256.0 * n) / 8192.0;
    realA[2 * n]     = q15_floor(32768.0 * 0.5 * (1.0 - sin(th)) + 0.5);
    realA[2 * n + 1] = q15_floor(32768.0 * 0.5 * (-cos(th)) + 0.5);
    realB[2 * n]     = q15_floor(32768.0 * 0.5 * (1.0 + sin(th)) + 0.5);   /* saturates to 32767 at n = 8 */
    realB[2 * n + 1] = q15_floor(32768.0 * 0.5 * (cos(th)) + 0.5);
  }
  gen_mel();
  inited = 1;
  return 0;
}

const int16_t *eo_table_twiddle1024(void) { eo_init(); return tw1024; }
const int16_t *eo_table_twiddle16(void) { eo_init(); return tw16; }
const int16_t *eo_table_realA(void) { eo_init(); return realA; }
const int16_t *eo_table_realB(void) { eo_init(); return realB; }
const int16_t *eo_table_mel(int *n, const int16_t **starts, const int16_t **counts)
{
  eo_init();
  if (n) *n = mel_n;
  if (starts) *starts = mel_start;
  if (counts) *counts = mel_cnt;
  return mel_w;
}

/* ------------------------------------------------------------------------------------------
 * Radix-4 DIF complex q15 FFT, SIMD branch: arm_cfft_radix4_q15.c:147-565
 * (arm_radix4_butterfly_q15) followed by arm_bitreversal_16 (arm_bitreversal2.c:110-134),
 * as dispatched by arm_cfft_q15 (arm_cfft_q15.c:695-746).
 * ---------------------------------------------------------------------------------------- */
static void radix4_q15(cq15 *x, int n, const int16_t *tw)
{
  int n2 = n >> 2, n1;
  int mod = 1;
  /* first stage (:192-330): inputs >>2 (two SHADD16 with 0), saturating adds, halved sum */
  for (int j = 0; j < n2; j++) {
    cq15 a = x[j], b = x[j + n2], c = x[j + 2 * n2], d = x[j + 3 * n2];
    a.re >>= 2; a.im >>= 2; b.re >>= 2; b.im >>= 2;
    c.re >>= 2; c.im >>= 2; d.re >>= 2; d.im >>= 2;
    cq15 R = { qadd(a.re, c.re), qadd(a.im, c.im) };
    cq15 S = { qsub(a.re, c.re), qsub(a.im, c.im) };
    cq15 T = { qadd(b.re, d.re), qadd(b.im, d.im) };
    x[j].re = hadd(R.re, T.re);
    x[j].im = hadd(R.im, T.im);
    cq15 R2 = { qsub(R.re, T.re), qsub(R.im, T.im) };
    x[j + n2] = twiddle(tw[4 * j], tw[4 * j + 1], R2);                  /* W^(2j)  :249-271 */
    cq15 U = { qsub(b.re, d.re), qsub(b.im, d.im) };
    cq15 Ra = { qsub(S.re, U.im), qadd(S.im, U.re) };                   /* QASX :285 */
    cq15 Sa = { qadd(S.re, U.im), qsub(S.im, U.re) };                   /* QSAX :287 */
    x[j + 2 * n2] = twiddle(tw[2 * j], tw[2 * j + 1], Sa);              /* W^j     :296-311 */
    x[j + 3 * n2] = twiddle(tw[6 * j], tw[6 * j + 1], Ra);              /* W^(3j)  :313-327 */
  }
  /* middle stages (:333-479) */
  mod <<= 2;
  for (int k = n / 4; k > 4; k >>= 2) {
    n1 = n2;
    n2 >>= 2;
    for (int j = 0; j < n2; j++) {
      int ic = j * mod;
      int32_t c1 = tw[2 * ic], s1 = tw[2 * ic + 1];
      int32_t c2 = tw[4 * ic], s2 = tw[4 * ic + 1];
      int32_t c3 = tw[6 * ic], s3 = tw[6 * ic + 1];
      for (int i0 = j; i0 < n; i0 += n1) {
        cq15 a = x[i0], b = x[i0 + n2], c = x[i0 + 2 * n2], d = x[i0 + 3 * n2];
        cq15 R = { qadd(a.re, c.re), qadd(a.im, c.im) };
        cq15 S = { qsub(a.re, c.re), qsub(a.im, c.im) };
        cq15 T = { qadd(b.re, d.re), qadd(b.im, d.im) };
        x[i0].re = hadd(R.re, T.re) >> 1;                               /* SHADD16 twice :372-374 */
        x[i0].im = hadd(R.im, T.im) >> 1;
        cq15 R2 = { hsub(R.re, T.re), hsub(R.im, T.im) };               /* SHSUB16 :378 */
        x[i0 + n2] = twiddle(c2, s2, R2);
        cq15 U = { qsub(b.re, d.re), qsub(b.im, d.im) };
        cq15 Ra = { hsub(S.re, U.im), hadd(S.im, U.re) };               /* SHASX :412 */
        cq15 Sa = { hadd(S.re, U.im), hsub(S.im, U.re) };               /* SHSAX :415 */
        x[i0 + 2 * n2] = twiddle(c1, s1, Sa);
        x[i0 + 3 * n2] = twiddle(c3, s3, Ra);
      }
    }
    mod <<= 2;
  }
  /* last stage (:492-557): no twiddles */
  for (int j = 0; j < n / 4; j++) {
    cq15 a = x[4 * j], b = x[4 * j + 1], c = x[4 * j + 2], d = x[4 * j + 3];
    cq15 R = { qadd(a.re, c.re), qadd(a.im, c.im) };
    cq15 T = { qadd(b.re, d.re), qadd(b.im, d.im) };
    cq15 S = { qsub(a.re, c.re), qsub(a.im, c.im) };
    cq15 U = { qsub(b.re, d.re), qsub(b.im, d.im) };
    x[4 * j].re = hadd(R.re, T.re);     x[4 * j].im = hadd(R.im, T.im);
    x[4 * j + 1].re = hsub(R.re, T.re); x[4 * j + 1].im = hsub(R.im, T.im);
    x[4 * j + 2].re = hadd(S.re, U.im); x[4 * j + 2].im = hsub(S.im, U.re);   /* SHSAX :540 */
    x[4 * j + 3].re = hsub(S.re, U.im); x[4 * j + 3].im = hadd(S.im, U.re);   /* SHASX :544 */
  }
}

static void bitreverse(cq15 *x, int n, int bits)
{
  /* armBitRevIndexTable_fixed_N lists the swap pairs of a plain bit reversal */
  for (int i = 0; i < n; i++) {
    int r = 0;
    for (int b = 0; b < bits; b++) r |= ((i >> b) & 1) << (bits - 1 - b);
    if (r > i) { cq15 t = x[i]; x[i] = x[r]; x[r] = t; }
  }
}

/* ------------------------------------------------------------------------------------------
 * arm_sqrt_q31: lib/CMSIS/DSP/Source/FastMathFunctions/arm_sqrt_q31.c:50-135
 * ---------------------------------------------------------------------------------------- */
int32_t eo_sqrt_q31(int32_t in)
{
  if (in <= 0) return 0;
  int32_t number = in, signBits = __builtin_clz((uint32_t)number) - 1;
  int32_t shift = (signBits % 2) == 0 ? signBits : signBits - 1;
  number = (int32_t)((uint32_t)number << shift);
  int32_t half = number >> 1, temp1 = number;
  union { int32_t i; float f; } cv;
  volatile float tf = (float)number * 4.6566128731e-010f;     /* :82, single precision */
  cv.f = tf;
  cv.i = 0x5f3759df - (cv.i >> 1);                             /* :87 */
  volatile float tg = cv.f * 1073741824.0f;                    /* :92 */
  int32_t var1 = (int32_t)tg;
  for (int it = 0; it < 3; it++) {                              /* :95-112 */
    int32_t a = (int32_t)(((int64_t)var1 * var1) >> 31);
    int32_t b = (int32_t)(((int64_t)a * (int64_t)half) >> 31);
    int32_t c = (int32_t)(((int64_t)var1 * (int64_t)(0x30000000 - b)) >> 31);
    var1 = (int32_t)((uint32_t)c << 2);
  }
  var1 = (int32_t)((uint32_t)((int32_t)(((int64_t)temp1 * var1) >> 31)) << 1);   /* :115 */
  return var1 >> (shift / 2);                                                       /* :118-125 */
}

/* ------------------------------------------------------------------------------------------
 * dct2_q15 (audioprocessing.c:324-436) = reorder + arm_rfft_q15 (arm_rfft_q15.c:76-120:
 * 16-point CFFT + arm_split_rfft_q15 SIMD branch :264-325) + "take the real parts".
 * The twiddle multiply of a true DCT-II is skipped by the reference (:413-421).
 * ---------------------------------------------------------------------------------------- */
static void dct2_q15(const int16_t *mel, int16_t *out)
{
  int16_t v[32];
  for (int i = 0; i < 16; i++) { v[i] = mel[2 * i]; v[31 - i] = mel[2 * i + 1]; }   /* :349-382 */
  cq15 z[16];
  for (int i = 0; i < 16; i++) { z[i].re = v[2 * i]; z[i].im = v[2 * i + 1]; }
  radix4_q15(z, 16, tw16);
  bitreverse(z, 16, 4);
  int16_t re[33];
  for (int i = 1; i < 16; i++) {
    /* outR = (SMUSD(src[i], A[i]) + SMLAD(src[16-i], B[i])) >> 16            arm_rfft_q15.c:283-290 */
    uint32_t acc = (uint32_t)(z[i].re * realA[2 * i]) - (uint32_t)(z[i].im * realA[2 * i + 1]);
    acc += (uint32_t)(z[16 - i].re * realB[2 * i]) + (uint32_t)(z[16 - i].im * realB[2 * i + 1]);
    re[i] = (int16_t)(acc >> 16);
    re[32 - i] = re[i];                                                        /* conjugate half :305-307 */
  }
  re[16] = (int16_t)((z[0].re - z[0].im) >> 1);                               /* :315 */
  re[0]  = (int16_t)((z[0].re + z[0].im) >> 1);                               /* :318 */
  for (int k = 0; k < 32; k++) out[k] = re[k];                                 /* audioprocessing.c:426-434 */
}

/* ------------------------------------------------------------------------------------------
 * audioCalcMFCCs: firmware/src/audioprocessing.c:116-215
 * ---------------------------------------------------------------------------------------- */
void eo_mfcc_frame_stages(const int16_t *pcm, int16_t *fft2048, int16_t *spect1024,
                          int16_t *mel32, int16_t *mfcc32)
{
  eo_init();
  static __thread cq15 x[1024];
  int16_t spect[1024], mel[32];
  for (int i = 0; i < 1024; i++) { x[i].re = pcm[i]; x[i].im = 0; }           /* :135-138 */
  radix4_q15(x, 1024, tw1024);                                                /* :139 */
  bitreverse(x, 1024, 10);
  if (fft2048) for (int i = 0; i < 1024; i++) { fft2048[2 * i] = (int16_t)x[i].re; fft2048[2 * i + 1] = (int16_t)x[i].im; }
  for (int i = 0; i < 1024; i++) {                                             /* cmpl_mag_q15 :294-307 */
    int32_t sum = (int32_t)((uint32_t)(x[i].re * x[i].re) + (uint32_t)(x[i].im * x[i].im));
    spect[i] = (int16_t)(eo_sqrt_q31(sum) >> 16);
  }
  if (spect1024) memcpy(spect1024, spect, sizeof spect);
  const int16_t *w = mel_w;
  for (int m = 0; m < 32; m++) {                                               /* :157-170 */
    int32_t acc = 0;
    for (int f = mel_start[m]; f < mel_start[m] + mel_cnt[m]; f++) acc += spect[f] * (*w++);
    mel[m] = (int16_t)(acc / 128);
  }
  if (mel32) memcpy(mel32, mel, sizeof mel);
  dct2_q15(mel, mfcc32);                                                       /* :197-205 */
}

void eo_mfcc_frame(const int16_t *pcm, int16_t *mfcc32) { eo_mfcc_frame_stages(pcm, 0, 0, 0, mfcc32); }

void eo_mfcc_batch(const int16_t *pcm, size_t n_frames, int16_t *mfcc)
{
  for (size_t f = 0; f < n_frames; f++) eo_mfcc_frame(pcm + f * 1024, mfcc + f * 32);
}

/* mfccToNetInput, NNoM branch: firmware/src/app.c:684-693; NNOM_INPUT_* weights.h:162-164 */
static int32_t in_scale = 1, in_min = -128, in_max = 127;
void eo_net_input_row(const int16_t *mfcc32, int8_t *row13)
{
  for (int c = 0; c < EO_N_COEFF; c++) {
    int16_t t = (int16_t)(mfcc32[c] / in_scale);
    t = t > in_max ? (int16_t)in_max : t;
    t = t < in_min ? (int16_t)in_min : t;
    row13[c] = (int8_t)t;
  }
}

/* ------------------------------------------------------------------------------------------
 * int8 NNoM net
 * ---------------------------------------------------------------------------------------- */
typedef struct {
  uint32_t kind, cout, kh, kw, cin, rshift, lshift, relu, pool;
  const int8_t *w, *b;
} eo_layer;
static eo_layer layers[16];
static uint32_t n_layers, net_h, net_w, net_c;
static uint8_t *blob;

int eo_nnom_load(const char *path)
{
  FILE *fh = fopen(path, "rb");
  if (!fh) return -1;
  fseek(fh, 0, SEEK_END);
  long sz = ftell(fh);
  fseek(fh, 0, SEEK_SET);
  free(blob);
  blob = (uint8_t *)malloc((size_t)sz);
  if (fread(blob, 1, (size_t)sz, fh) != (size_t)sz) { fclose(fh); return -2; }
  fclose(fh);
  if (memcmp(blob, "EDNNQ7\0\0", 8)) return -3;
  const uint8_t *p = blob + 8;
  uint32_t ver;
  memcpy(&ver, p, 4); memcpy(&n_layers, p + 4, 4); p += 8;
  if (ver != 1 || n_layers > 16) return -4;
  for (uint32_t i = 0; i < n_layers; i++) {
    uint32_t h[9];
    memcpy(h, p, sizeof h); p += sizeof h;
    eo_layer *L = &layers[i];
    L->kind = h[0]; L->cout = h[1]; L->kh = h[2]; L->kw = h[3]; L->cin = h[4];
    L->rshift = h[5]; L->lshift = h[6]; L->relu = h[7]; L->pool = h[8];
    L->w = (const int8_t *)p; p += (size_t)L->cout * L->kh * L->kw * L->cin;
    L->b = (const int8_t *)p; p += L->cout;
    p += (size_t)(-(p - blob)) & 3;
  }
  int32_t t[3]; uint32_t s[3];
  memcpy(t, p, 12); memcpy(s, p + 12, 12);
  in_scale = t[0]; in_min = t[1]; in_max = t[2];
  net_h = s[0]; net_w = s[1]; net_c = s[2];
  return 0;
}

static inline int32_t ssat8(int32_t v) { return v > 127 ? 127 : (v < -128 ? -128 : v); }
static inline int32_t usat(int32_t v, int bits) { int32_t m = (1 << bits) - 1; return v < 0 ? 0 : (v > m ? m : v); }

/* valid, stride-1 HWC convolution: nnom_local.c:425-484 == arm_convolve_HWC_q7_{basic,fast}_nonsquare
 * (CMSIS-NN; same integer result), followed by the ReLU tail nnom_activation.c:118-126 */
static void conv_q7(const eo_layer *L, const int8_t *in, int H, int W, int8_t *out)
{
  int Ho = H - (int)L->kh + 1, Wo = W - (int)L->kw + 1, C = (int)L->cin, O = (int)L->cout;
  for (int y = 0; y < Ho; y++)
    for (int x = 0; x < Wo; x++)
      for (int o = 0; o < O; o++) {
        int32_t acc = (int32_t)((uint32_t)(int32_t)L->b[o] << L->lshift) + (1 << (L->rshift - 1));
        for (int ky = 0; ky < (int)L->kh; ky++)
          for (int kx = 0; kx < (int)L->kw; kx++)
            for (int c = 0; c < C; c++)
              acc += in[((y + ky) * W + (x + kx)) * C + c] * L->w[((o * (int)L->kh + ky) * (int)L->kw + kx) * C + c];
        int32_t v = ssat8(acc >> L->rshift);
        if (L->relu && v < 0) v = 0;
        out[(y * Wo + x) * O + o] = (int8_t)v;
      }
}

/* (p x 1) max pool, stride (p, 1), valid: nnom_maxpool.c:109-152 -> local_maxpool_q7_HWC nnom_local.c:117-160 */
static void maxpool_rows(const int8_t *in, int H, int W, int C, int p, int8_t *out)
{
  int Ho = (H - p) / p + 1;
  for (int y = 0; y < Ho; y++)
    for (int i = 0; i < W * C; i++) {
      int m = -129;
      for (int k = 0; k < p; k++) { int v = in[(y * p + k) * W * C + i]; if (v > m) m = v; }
      out[y * W * C + i] = (int8_t)m;
    }
}

/* arm_softmax_q7 (CMSIS-NN SoftmaxFunctions/arm_softmax_q7.c: reference branch :217-254, SIMD branch
 * :83-215) and NNoM's local_softmax_q7 (nnom_local.c:948-1003) */
void eo_softmax_q7(const int8_t *in, int n, int8_t *out, int mode)
{
  if (mode == EO_SOFTMAX_NNOM_LOCAL) {
    int base = -257;
    for (int i = 0; i < n; i++) if (in[i] > base) base = in[i];
    base -= 8;
    int32_t sum = 0;
    for (int i = 0; i < n; i++) if (in[i] > base) sum += 1 << (uint8_t)usat(in[i] - base, 5);
    int ob = 0x100000 / sum;
    for (int i = 0; i < n; i++)
      out[i] = in[i] > base ? (int8_t)ssat8(ob >> (uint8_t)usat(13 + base - in[i], 5)) : 0;
    return;
  }
  int base = -128;
  for (int i = 0; i < n; i++) if (in[i] > base) base = in[i];
  base -= 8;
  int32_t sum = 0;
  int nv = (mode == EO_SOFTMAX_CMSIS_SIMD) ? (n & ~3) : 0;   /* elements handled 4 at a time with 8-bit wrap */
  for (int i = 0; i < n; i++) {
    int d = in[i] - base;
    if (i < nv) d = (int8_t)(uint8_t)d;                      /* __SSUB8 wraps modulo 256 (:127) */
    sum += 1 << usat(d, 3);
  }
  int ob = (1 << 20) / sum;
  for (int i = 0; i < n; i++) {
    int d;
    if (i < nv) {
      int t = (int8_t)(uint8_t)(13 - in[i]);                 /* __SSUB8(pad, in) :177 */
      d = (int8_t)(uint8_t)(t + (int8_t)(uint8_t)base);      /* __SADD8(in, baseV) :178 */
    } else {
      d = 13 + base - in[i];
    }
    out[i] = (int8_t)ssat8(ob >> usat(d, 5));
  }
}

static int forward(const int8_t *in403, int8_t *logits, int8_t *softmax, int mode, int8_t *taps)
{
  if (!n_layers) return -1;
  static __thread int8_t bufA[65536], bufB[65536];
  int H = (int)net_h, W = (int)net_w, C = (int)net_c, tap = 0;
  memcpy(bufA, in403, (size_t)(H * W * C));
  int8_t *cur = bufA, *nxt = bufB, *t;
  for (uint32_t i = 0; i < n_layers; i++) {
    const eo_layer *L = &layers[i];
    if (L->kind == 1) {
      conv_q7(L, cur, H, W, nxt);
      H = H - (int)L->kh + 1; W = W - (int)L->kw + 1; C = (int)L->cout;
      t = cur; cur = nxt; nxt = t;
      if (taps) memcpy(taps + 4096 * tap++, cur, (size_t)(H * W * C));
      if (L->pool) {
        maxpool_rows(cur, H, W, C, (int)L->pool, nxt);
        H = (H - (int)L->pool) / (int)L->pool + 1;
        t = cur; cur = nxt; nxt = t;
        if (taps) memcpy(taps + 4096 * tap++, cur, (size_t)(H * W * C));
      }
    } else {
      /* dense: nnom_dense.c:92-119 -> arm_fully_connected_q7_opt / local_fully_connected_q7_opt
       * (nnom_local.c:804-921); weights de-interleaved by tools/convert_nnom_weights.py */
      int n_in = H * W * C;
      if ((uint32_t)n_in != L->cin) return -2;
      for (uint32_t r = 0; r < L->cout; r++) {
        int32_t acc = (int32_t)((uint32_t)(int32_t)L->b[r] << L->lshift) + (1 << (L->rshift - 1));
        for (int j = 0; j < n_in; j++) acc += cur[j] * L->w[r * (uint32_t)n_in + (uint32_t)j];
        int32_t v = ssat8(acc >> L->rshift);
        if (L->relu && v < 0) v = 0;
        nxt[r] = (int8_t)v;
      }
      H = 1; W = 1; C = (int)L->cout;
      t = cur; cur = nxt; nxt = t;
      if (taps) memcpy(taps + 4096 * tap++, cur, (size_t)C);
    }
  }
  const int n_out = H * W * C;      /* 10 for the reference model */
  if (n_out > 256) return -3;
  if (logits) memcpy(logits, cur, (size_t)n_out);
  if (softmax || taps) {
    int8_t sm[256];
    eo_softmax_q7(cur, n_out, sm, mode);
    if (softmax) memcpy(softmax, sm, (size_t)n_out);
    if (taps) memcpy(taps + 4096 * tap++, sm, (size_t)n_out);
  }
  return tap;
}

int eo_nnom_forward(const int8_t *in403, int8_t *logits10, int8_t *softmax10, int mode)
{
  int r = forward(in403, logits10, softmax10, mode, 0);
  return r < 0 ? r : 0;
}

int eo_nnom_forward_batch(const int8_t *in, size_t n, int8_t *logits, int8_t *softmax, int mode)
{
  for (size_t i = 0; i < n; i++) {
    int r = forward(in + i * EO_NET_IN, logits ? logits + i * 10 : 0, softmax ? softmax + i * 10 : 0, mode, 0);
    if (r < 0) return r;
  }
  return 0;
}

/* any model of the blob's layer family (tools/convert_nnom_weights.py): in [n][in_h*in_w*in_c], outputs [n][n_out] */
int eo_nnom_forward_generic(const int8_t *in, size_t n, int8_t *logits, int8_t *softmax, int mode)
{
  if (!n_layers) return -1;
  const size_t in_size = (size_t)net_h * net_w * net_c;
  const eo_layer *last = &layers[n_layers - 1];
  if (last->kind != 2) return -4;
  const size_t n_out = last->cout;
  for (size_t i = 0; i < n; i++) {
    int r = forward(in + i * in_size, logits ? logits + i * n_out : 0, softmax ? softmax + i * n_out : 0, mode, 0);
    if (r < 0) return r;
  }
  return 0;
}
int eo_nnom_in_size(void) { return (int)(net_h * net_w * net_c); }
int eo_nnom_out_size(void) { return n_layers ? (int)layers[n_layers - 1].cout : 0; }

int eo_nnom_forward_layers(const int8_t *in403, int8_t *buf8x4096)
{
  return forward(in403, 0, 0, EO_SOFTMAX_CMSIS_REF, buf8x4096);
}

/* appHifMfccAndInference: firmware/src/app.c:167-221 over a batch of clips */
int eo_kws_batch(const int16_t *pcm, size_t n_clips, size_t clip_stride, int8_t *logits,
                 int8_t *softmax, int8_t *net_in, int mode)
{
  int8_t in[EO_NET_IN];
  int16_t mf[32];
  for (size_t c = 0; c < n_clips; c++) {
    for (int f = 0; f < EO_N_FRAMES; f++) {
      eo_mfcc_frame(pcm + c * clip_stride + (size_t)f * 1024, mf);
      eo_net_input_row(mf, in + f * EO_N_COEFF);
    }
    if (net_in) memcpy(net_in + c * EO_NET_IN, in, EO_NET_IN);
    int r = forward(in, logits ? logits + c * 10 : 0, softmax ? softmax + c * 10 : 0, mode, 0);
    if (r < 0) return r;
  }
  return 0;
}
