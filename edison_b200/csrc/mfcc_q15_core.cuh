// mfcc_q15_core.cuh -- the q15 MFCC front-end (audioCalcMFCCs, firmware/src/audioprocessing.c:116-215)
// laid out for one warp per 1024-sample frame.
//
// The 1024-point radix-4 DIF FFT (arm_cfft_q15 -> arm_radix4_butterfly_q15, SIMD branch,
// arm_cfft_radix4_q15.c:147-565) has five stages; index bits of position p are b9..b0 and stage s
// combines the four positions that differ in bits (11-2s, 10-2s).  A lane keeps 32 complex points
// in registers, i.e. it owns five index bits:
//
//   phase A  lane owns b9 b8 b7 b6 b0   p = 64 q + 2 L + e          stages 1, 2   (input is real)
//   phase B  lane owns b5 b4 b3 b2 b0   p = 64 (L>>1) + 4 t + 2 (L&1) + e   stages 3, 4
//   phase C  lane owns b4..b0 of 8 butterflies  p = 32 L + 4 i + {0..3}      stage 5 (only the
//            two outputs per butterfly that land in bins 0..511), magnitude
//
// with two exchanges through shared memory (int32 planes re[], im[], padded 4 words per 64).
// Bit reversal (arm_bitreversal_16) is folded into phase C's addressing: position p holds bin
// rev10(p).  Then mel (32 sparse dot products), and, batched over the 8 frames of a group,
// dct2_q15 (audioprocessing.c:324-436) with 4 lanes per frame.
//
// Every function is lane-local between __syncwarp() points, so the same source runs under the
// host emulation harness (tests/host_emu) that checks the index arithmetic against the oracle
// without a GPU.
#pragma once
#include "q15_arith.cuh"

namespace edison {

constexpr int kFrameLen = 1024;
constexpr int kNMel = 32;
constexpr int kNCoeff = 13;       // aiGetInputShape, firmware/src/ai/ai.c:205-216
constexpr int kGroup = 8;         // frames per warp iteration (DCT batch)
constexpr int kMelMaxW = 1024;    // capacity for melMtxCompact (915 in the reference header)

// exchange buffer: two planes of 1024 int32 padded by 4 words every 64
constexpr int kXchPlane = 1024 + 4 * 16;
ED_HD int xaddr(int p) { return p + ((p >> 6) << 2); }

// ---- constant tables, one copy per CTA in shared memory -----------------------------------------
struct MfccTables {
  I2 tw1[3][256];   // stage 1: [0] W^j  [1] W^2j  [2] W^3j          (co, si) of twiddleCoef_1024_q15
  I2 tw2[3][64];    // stage 2: W^(4 j (k+1))
  I2 tw3[3][16];    // stage 3: W^(16 j (k+1))
  I2 tw4[3][4];     // stage 4: W^(64 j (k+1))
  I2 tw16[12];      // twiddleCoef_16_q15 (DCT's 16-point CFFT)
  I2 realA[16];     // realCoefAQ15 at stride 512 (arm_rfft_init_q15.c:213-222, fftLenReal 32)
  I2 realB[16];
  uint32_t sqrt_exc[kSqrtExcWords];
  int16_t mel_w[kMelMaxW];      // melMtxCompact   (firmware/src/audio/mel_constants.h:2586)
  int16_t mel_start[kNMel];     // melCompFStarts
  int16_t mel_cnt[kNMel];       // melCompFCount
  int16_t mel_off[kNMel];       // prefix sum of counts
  int in_scale, in_min, in_max; // NNOM_INPUT_SCALE/MIN/MAX (weights.h:162-164)
  int pad_;
};

// ---- phase A: load, stages 1 and 2 ---------------------------------------------------------------
// w[q] = the 32-bit word holding samples 64q+2L (low half) and 64q+2L+1 (high half).
// Returns max(|min|, |max|) proxy: (max sample, min sample) folded into one "peak" value.
ED_HD int phase_a(int lane, const uint32_t w[16], const MfccTables &T, int *xr, int *xi) {
  int re[32], im[32];
  int mx = -32768, mn = 32767;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int q = 0; q < 16; q++) {
    int lo = (int)(int16_t)(w[q] & 0xffffu), hi = (int)w[q] >> 16;
    re[2 * q] = lo; re[2 * q + 1] = hi;
    mx = imax(mx, imax(lo, hi));
    mn = imin(mn, imin(lo, hi));
  }
  // stage 1 (first-stage formulas, imaginary input zero): butterflies j = 64 qp + 2L + e
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int qp = 0; qp < 4; qp++) {
    const I4 w1 = *reinterpret_cast<const I4 *>(&T.tw1[0][64 * qp + 2 * lane]);
    const I4 w2 = *reinterpret_cast<const I4 *>(&T.tw1[1][64 * qp + 2 * lane]);
    const I4 w3 = *reinterpret_cast<const I4 *>(&T.tw1[2][64 * qp + 2 * lane]);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int e = 0; e < 2; e++) {
      const int c1 = e ? w1.z : w1.x, s1 = e ? w1.w : w1.y;
      const int c2 = e ? w2.z : w2.x, s2 = e ? w2.w : w2.y;
      const int c3 = e ? w3.z : w3.x, s3 = e ? w3.w : w3.y;
      const int a = re[2 * qp + e] >> 2, b = re[2 * (qp + 4) + e] >> 2;
      const int c = re[2 * (qp + 8) + e] >> 2, d = re[2 * (qp + 12) + e] >> 2;
      const int R = a + c, S = a - c, Tt = b + d, U = b - d;
      re[2 * qp + e] = (R + Tt) >> 1;
      im[2 * qp + e] = 0;
      const int r2 = R - Tt;
      re[2 * (qp + 4) + e] = (c2 * r2) >> 16;
      im[2 * (qp + 4) + e] = (-s2 * r2) >> 16;
      re[2 * (qp + 8) + e] = (c1 * S - s1 * U) >> 16;       // Sa = (S, -U)
      im[2 * (qp + 8) + e] = (-c1 * U - s1 * S) >> 16;
      re[2 * (qp + 12) + e] = (c3 * S + s3 * U) >> 16;      // Ra = (S, U)
      im[2 * (qp + 12) + e] = (c3 * U - s3 * S) >> 16;
    }
  }
  // stage 2 (middle-stage formulas; operands are bounded by 16384 so no saturation)
  {
    const I4 w1 = *reinterpret_cast<const I4 *>(&T.tw2[0][2 * lane]);
    const I4 w2 = *reinterpret_cast<const I4 *>(&T.tw2[1][2 * lane]);
    const I4 w3 = *reinterpret_cast<const I4 *>(&T.tw2[2][2 * lane]);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int g = 0; g < 4; g++) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
      for (int e = 0; e < 2; e++) {
        const I2 t1 = e ? I2{w1.z, w1.w} : I2{w1.x, w1.y};
        const I2 t2 = e ? I2{w2.z, w2.w} : I2{w2.x, w2.y};
        const I2 t3 = e ? I2{w3.z, w3.w} : I2{w3.x, w3.y};
        const int q0 = 2 * (4 * g) + e;
        bfly_mid<false>(re[q0], im[q0], re[q0 + 2], im[q0 + 2], re[q0 + 4], im[q0 + 4], re[q0 + 6], im[q0 + 6],
                        t1, t2, t3);
      }
    }
  }
  // exchange 1: position p = 64 q + 2 L + e
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int q = 0; q < 16; q++) {
    const int a = xaddr(64 * q + 2 * lane);
    *reinterpret_cast<I2 *>(xr + a) = I2{re[2 * q], re[2 * q + 1]};
    *reinterpret_cast<I2 *>(xi + a) = I2{im[2 * q], im[2 * q + 1]};
  }
  return imax(mx, -mn - 1);  // >= |x| - 1 for every sample; compared against kNoSatPeak
}

// A frame whose peak is at most kNoSatPeak cannot saturate any QADD16/QSUB16 of stages 3..5:
// the complex magnitude after stage k is bounded by peak/2 + 3 + 2.2 (k-1), so sums of
// two components stay below 32767.
constexpr int kNoSatPeak = 32000;

// ---- phase B: stages 3 and 4 ---------------------------------------------------------------------
template <bool SAT>
ED_HD void phase_b(int lane, const MfccTables &T, int *xr, int *xi) {
  int re[32], im[32];  // index 2 t + e, t = bits 5..2 of p
  const int base = xaddr(64 * (lane >> 1) + 2 * (lane & 1));
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int t = 0; t < 16; t++) {
    const I2 a = *reinterpret_cast<const I2 *>(xr + base + 4 * t);
    const I2 b = *reinterpret_cast<const I2 *>(xi + base + 4 * t);
    re[2 * t] = a.x; re[2 * t + 1] = a.y;
    im[2 * t] = b.x; im[2 * t + 1] = b.y;
  }
  // stage 3: n2 = 16, twiddle index j3 = p & 15 = 4 tl + 2 (L&1) + e
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int tl = 0; tl < 4; tl++) {
    const int j = 4 * tl + 2 * (lane & 1);
    const I4 w1 = *reinterpret_cast<const I4 *>(&T.tw3[0][j]);
    const I4 w2 = *reinterpret_cast<const I4 *>(&T.tw3[1][j]);
    const I4 w3 = *reinterpret_cast<const I4 *>(&T.tw3[2][j]);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int e = 0; e < 2; e++) {
      const I2 t1 = e ? I2{w1.z, w1.w} : I2{w1.x, w1.y};
      const I2 t2 = e ? I2{w2.z, w2.w} : I2{w2.x, w2.y};
      const I2 t3 = e ? I2{w3.z, w3.w} : I2{w3.x, w3.y};
      const int i0 = 2 * tl + e;  // th = 0; th adds 8 to the local index
      bfly_mid<SAT>(re[i0], im[i0], re[i0 + 8], im[i0 + 8], re[i0 + 16], im[i0 + 16], re[i0 + 24], im[i0 + 24],
                    t1, t2, t3);
    }
  }
  // stage 4: n2 = 4, twiddle index j4 = p & 3 = 2 (L&1) + e
  {
    const int j = 2 * (lane & 1);
    const I4 w1 = *reinterpret_cast<const I4 *>(&T.tw4[0][j]);
    const I4 w2 = *reinterpret_cast<const I4 *>(&T.tw4[1][j]);
    const I4 w3 = *reinterpret_cast<const I4 *>(&T.tw4[2][j]);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int th = 0; th < 4; th++) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
      for (int e = 0; e < 2; e++) {
        const I2 t1 = e ? I2{w1.z, w1.w} : I2{w1.x, w1.y};
        const I2 t2 = e ? I2{w2.z, w2.w} : I2{w2.x, w2.y};
        const I2 t3 = e ? I2{w3.z, w3.w} : I2{w3.x, w3.y};
        const int i0 = 8 * th + e;  // tl adds 2 to the local index
        bfly_mid<SAT>(re[i0], im[i0], re[i0 + 2], im[i0 + 2], re[i0 + 4], im[i0 + 4], re[i0 + 6], im[i0 + 6],
                      t1, t2, t3);
      }
    }
  }
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int t = 0; t < 16; t++) {
    *reinterpret_cast<I2 *>(xr + base + 4 * t) = I2{re[2 * t], re[2 * t + 1]};
    *reinterpret_cast<I2 *>(xi + base + 4 * t) = I2{im[2 * t], im[2 * t + 1]};
  }
}

ED_HD int rev3(int i) { return ((i & 1) << 2) | (i & 2) | ((i >> 2) & 1); }
ED_HD int rev5(int i) {
  return ((i & 1) << 4) | ((i & 2) << 2) | (i & 4) | ((i >> 2) & 2) | ((i >> 4) & 1);
}

// ---- phase C: stage 5 (outputs 0 and 2 of each butterfly), cmpl_mag_q15 --------------------------
// Butterfly j = 8 L + i sits at positions 4j..4j+3; its outputs 4j and 4j+2 are bins
// r = rev8(j) = 32 rev3(i) + rev5(L) and r + 256.  Outputs 4j+1, 4j+3 are bins >= 512, which
// nothing reads (the mel filters end at bin 486; audioprocessing.c:157-170).
template <bool SAT>
ED_HD void phase_c(int lane, const MfccTables &T, const int *xr, const int *xi, int *spect) {
  const int rl = rev5(lane);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < 8; i++) {
    const int a = xaddr(32 * lane + 4 * i);
    const I4 vr = *reinterpret_cast<const I4 *>(xr + a);
    const I4 vi = *reinterpret_cast<const I4 *>(xi + a);
    const int Rr = qadd16<SAT>(vr.x, vr.z), Ri = qadd16<SAT>(vi.x, vi.z);
    const int Sr = qsub16<SAT>(vr.x, vr.z), Si = qsub16<SAT>(vi.x, vi.z);
    const int Tr = qadd16<SAT>(vr.y, vr.w), Ti = qadd16<SAT>(vi.y, vi.w);
    const int Ur = qsub16<SAT>(vr.y, vr.w), Ui = qsub16<SAT>(vi.y, vi.w);
    const int r = 32 * rev3(i) + rl;
    spect[r] = mag_q15((Rr + Tr) >> 1, (Ri + Ti) >> 1, T.sqrt_exc);
    spect[r + 256] = mag_q15((Sr + Ui) >> 1, (Si - Ur) >> 1, T.sqrt_exc);
  }
}

// ---- mel: lane m owns filter m (audioprocessing.c:157-170) ---------------------------------------
ED_HD int phase_mel(int lane, const MfccTables &T, const int *spect) {
  const int start = T.mel_start[lane], cnt = T.mel_cnt[lane], off = T.mel_off[lane];
  int acc = 0;
  for (int i = 0; i < cnt; i++) acc += spect[start + i] * (int)T.mel_w[off + i];
  return (int)(int16_t)(acc / 128);  // C '/' truncates toward zero; the cast wraps like (q15_t)
}

// ---- dct2_q15 over the 8 frames of a group, 4 lanes per frame ------------------------------------
// mel: [8][32] ints; z: [8][16] (re, im); out: [8][32] ints (q15 values)
ED_HD int dct_src(const int *mel, int n2) {  // v[n2] after the reorder :349-382
  return n2 < 16 ? mel[2 * n2] : mel[2 * (31 - n2) + 1];
}

ED_HD void dct_stage1(int lane, const MfccTables &T, const int *mel, I2 *z) {
  const int fr = lane >> 2, j = lane & 3;
  const int *m = mel + 32 * fr;
  I2 *zz = z + 16 * fr;
  int ar = dct_src(m, 2 * j), ai = dct_src(m, 2 * j + 1);
  int br = dct_src(m, 2 * (j + 4)), bi = dct_src(m, 2 * (j + 4) + 1);
  int cr = dct_src(m, 2 * (j + 8)), ci = dct_src(m, 2 * (j + 8) + 1);
  int dr = dct_src(m, 2 * (j + 12)), di = dct_src(m, 2 * (j + 12) + 1);
  bfly_first(ar, ai, br, bi, cr, ci, dr, di, T.tw16[j], T.tw16[2 * j], T.tw16[3 * j]);
  zz[j] = I2{ar, ai};
  zz[j + 4] = I2{br, bi};
  zz[j + 8] = I2{cr, ci};
  zz[j + 12] = I2{dr, di};
}

ED_HD void dct_stage2(int lane, I2 *z) {
  const int fr = lane >> 2, j = lane & 3;
  I2 *zz = z + 16 * fr + 4 * j;
  I2 a = zz[0], b = zz[1], c = zz[2], d = zz[3];
  bfly_last<true>(a.x, a.y, b.x, b.y, c.x, c.y, d.x, d.y);
  zz[0] = a; zz[1] = b; zz[2] = c; zz[3] = d;
}

ED_HD int rev4(int i) { return ((i & 1) << 3) | ((i & 2) << 1) | ((i >> 1) & 2) | ((i >> 3) & 1); }

// arm_split_rfft_q15 (arm_rfft_q15.c:241-382, SIMD branch :264-325) + "take the real parts"
// (audioprocessing.c:426-434).  Task (fr, i): i in 0..15; i == 0 also produces coefficient 16.
ED_HD void dct_split(int task, const MfccTables &T, const I2 *z, int *out) {
  const int fr = task >> 4, i = task & 15;
  const I2 *zz = z + 16 * fr;
  int *o = out + 32 * fr;
  if (i == 0) {
    const I2 z0 = zz[0];
    o[0] = (int)(int16_t)((z0.x + z0.y) >> 1);   // :318
    o[16] = (int)(int16_t)((z0.x - z0.y) >> 1);  // :315
  } else {
    const I2 p = zz[rev4(i)], q = zz[rev4(16 - i)];
    const I2 A = T.realA[i], B = T.realB[i];
    const unsigned acc = (unsigned)(p.x * A.x) - (unsigned)(p.y * A.y) + (unsigned)(q.x * B.x) + (unsigned)(q.y * B.y);
    const int v = (int)(int16_t)(acc >> 16);     // :283-290
    o[i] = v;
    o[32 - i] = v;                               // conjugate half :305-307
  }
}

// mfccToNetInput, NNoM branch (firmware/src/app.c:684-693)
ED_HD int net_input_q7(int mfcc, const MfccTables &T) {
  int t = (int)(int16_t)(mfcc / T.in_scale);
  t = t > T.in_max ? T.in_max : t;
  t = t < T.in_min ? T.in_min : t;
  return (int)(int8_t)t;
}

}  // namespace edison
