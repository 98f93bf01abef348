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
// with two exchanges through shared memory (int32 planes re[], im[]).  Phases A and B run as a
// two-trip loop over the parity e = b0 (16 points = one closed radix-4 x radix-4 block per trip),
// phase C as a loop over its 8 butterflies: the loops are NOT unrolled, which keeps the kernel's
// hot code inside the instruction cache (a fully unrolled version was 126 KB of SASS and spent
// most of its time waiting for instruction fetch; profiles/r01).
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

// exchange buffer: 1024 packed complex points (re in the low, im in the high 16 bits of one word; the
// FFT's intermediate values are q15 by construction).  Position p = 64 q + 2 l + e is stored at
//   e * 544 + 34 q + l          (q = p >> 6, l = (p >> 1) & 31, e = p & 1)
// so that every access pattern of the three phases is free of bank conflicts.
constexpr int kXchHalf = 16 * 34;
constexpr int kXchWords = 2 * kXchHalf;
ED_HD uint32_t pack_c(int re, int im) { return ((uint32_t)re & 0xffffu) | ((uint32_t)im << 16); }
ED_HD int c_re(uint32_t w) { return (int)(int16_t)(w & 0xffffu); }
ED_HD int c_im(uint32_t w) { return (int)w >> 16; }

// magnitude spectrum of the 4 frames of a half-group, as two byte planes (low / high byte of the q15
// magnitude) so that the mel filterbank can run on the u8 tensor cores: [plane][frame][bin], rows padded
// to 528 bytes (132 words) which makes the B-fragment loads conflict-free.
constexpr int kSub = 4;
constexpr int kSpectRowWords = 132;
constexpr int kSpectPlaneWords = kSub * kSpectRowWords;
constexpr int kSpectWords = 2 * kSpectPlaneWords;
constexpr int kMelPairsMax = 32;   // (16-filter tile, 32-bin k-step) pairs with a non-zero weight

// ---- constant tables, one copy per CTA in shared memory -----------------------------------------
struct MfccTables {
  I2 tw1[3][2][4][32];  // stage 1: [k][e][qp][lane] = W^((k+1) j), j = 64 qp + 2 lane + e   (co, si)
  I2 tw2[3][2][32];     // stage 2: [k][e][lane]     = W^(4 (k+1) j), j = 2 lane + e
  I2 tw3[3][16];        // stage 3: W^(16 (k+1) j)
  I2 tw4[3][4];         // stage 4: W^(64 (k+1) j)
  I2 tw16[12];          // twiddleCoef_16_q15 (DCT's 16-point CFFT)
  I2 realA[16];         // realCoefAQ15 at stride 512 (arm_rfft_init_q15.c:213-222, fftLenReal 32)
  I2 realB[16];
  uint32_t sqrt_exc[kSqrtExcWords];
  // mel weights (melMtxCompact, firmware/src/audio/mel_constants.h:2586) as mma.m16n8k32 A fragments:
  // pair p covers filters 16 mt .. 16 mt + 15 and bins 32 ks .. 32 ks + 31; pairs with mt = 0 come first
  I4 mel_a[kMelPairsMax][32];
  uint8_t mel_ks[kMelPairsMax];
  int mel_n0, mel_n1;           // number of pairs with mt = 0 / mt = 1
  int in_scale, in_min, in_max; // NNOM_INPUT_SCALE/MIN/MAX (weights.h:162-164)
  int pad_;
};

// sample e (0 = low half, 1 = high half) of a packed pair, sign extended
ED_HD int unpack_half(uint32_t w, int e) {
#if defined(__CUDA_ARCH__)
  int r;  // PRMT, generic mode: selector nibble 8|n replicates the sign of byte n (__byte_perm masks that bit)
  asm("prmt.b32 %0, %1, 0, %2;" : "=r"(r) : "r"(w), "r"(e ? 0xBB32u : 0x9910u));
  return r;
#else
  return e ? (int)w >> 16 : (int)(int16_t)(w & 0xffffu);
#endif
}

// peak of the 32 samples a lane holds: >= |x| - 1 for every sample
ED_HD int lane_peak(const uint32_t w[16]) {
#if defined(__CUDA_ARCH__)
  uint32_t mx = w[0], mn = w[0];
#pragma unroll
  for (int q = 1; q < 16; q++) { mx = __vmaxs2(mx, w[q]); mn = __vmins2(mn, w[q]); }
  const int hi = max((int)mx >> 16, (int)(int16_t)(mx & 0xffffu));
  const int lo = min((int)mn >> 16, (int)(int16_t)(mn & 0xffffu));
  return max(hi, -lo - 1);
#else
  int hi = -32768, lo = 32767;
  for (int q = 0; q < 16; q++)
    for (int e = 0; e < 2; e++) { int v = unpack_half(w[q], e); hi = imax(hi, v); lo = imin(lo, v); }
  return imax(hi, -lo - 1);
#endif
}

// A frame whose peak is at most kNoSatPeak cannot saturate any QADD16/QSUB16 of stages 3..5:
// the complex magnitude after stage k is bounded by peak/2 + 3 + 2.2 (k-1), so sums of
// two components stay below 32767.
constexpr int kNoSatPeak = 32000;

// ---- phase A: stages 1 and 2 for the 16 points p = 64 q + 2 lane + e of one parity e ------------
// w[q] = the 32-bit word holding samples 64q+2L (low half) and 64q+2L+1 (high half).
ED_HD void phase_a(int lane, int e, const uint32_t w[16], const MfccTables &T, uint32_t *x) {
  int re[16], im[16];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int q = 0; q < 16; q++) re[q] = unpack_half(w[q], e) >> 2;  // first stage: inputs >> 2 (:200-212)
  // stage 1 (first-stage formulas, imaginary input zero): butterflies j = 64 qp + 2L + e
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int qp = 0; qp < 4; qp++) {
    const I2 w1 = T.tw1[0][e][qp][lane], w2 = T.tw1[1][e][qp][lane], w3 = T.tw1[2][e][qp][lane];
    const int a = re[qp], b = re[qp + 4], c = re[qp + 8], d = re[qp + 12];
    const int R = a + c, S = a - c, Tt = b + d, U = b - d;
    re[qp] = (R + Tt) >> 1;
    im[qp] = 0;
    const int r2 = R - Tt;
    re[qp + 4] = (w2.x * r2) >> 16;
    im[qp + 4] = (-w2.y * r2) >> 16;
    re[qp + 8] = (w1.x * S - w1.y * U) >> 16;    // Sa = (S, -U)
    im[qp + 8] = (-w1.x * U - w1.y * S) >> 16;
    re[qp + 12] = (w3.x * S + w3.y * U) >> 16;   // Ra = (S, U)
    im[qp + 12] = (w3.x * U - w3.y * S) >> 16;
  }
  // stage 2 (middle-stage formulas; operands are bounded by 16384 so no saturation)
  {
    const I2 w1 = T.tw2[0][e][lane], w2 = T.tw2[1][e][lane], w3 = T.tw2[2][e][lane];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int g = 0; g < 4; g++)
      bfly_mid_raw<false>(re[4 * g], im[4 * g], re[4 * g + 1], im[4 * g + 1], re[4 * g + 2], im[4 * g + 2], re[4 * g + 3],
                          im[4 * g + 3], w1, w2, w3);
  }
  // exchange 1 (q = 4g holds a finished value, the other three hold unshifted twiddle sums)
  const int base = e * kXchHalf + lane;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int q = 0; q < 16; q++) x[base + 34 * q] = (q & 3) ? pack_hi(re[q], im[q]) : pack_c(re[q], im[q]);
}

// ---- phase B: stages 3 and 4 for the 16 points p = 64 (L>>1) + 4 t + 2 (L&1) + e ----------------
template <bool SAT>
ED_HD void phase_b(int lane, int e, const MfccTables &T, uint32_t *x) {
  int re[16], im[16];  // index t = bits 5..2 of p
  const int base = e * kXchHalf + 34 * (lane >> 1) + (lane & 1);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int t = 0; t < 16; t++) {
    const uint32_t v = x[base + 2 * t];
    re[t] = c_re(v);
    im[t] = c_im(v);
  }
  // stage 3: n2 = 16, twiddle index j3 = p & 15 = 4 tl + 2 (L&1) + e
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int tl = 0; tl < 4; tl++) {
    const int j = 4 * tl + 2 * (lane & 1) + e;
    bfly_mid<SAT>(re[tl], im[tl], re[tl + 4], im[tl + 4], re[tl + 8], im[tl + 8], re[tl + 12], im[tl + 12],
                  T.tw3[0][j], T.tw3[1][j], T.tw3[2][j]);
  }
  // stage 4: n2 = 4, twiddle index j4 = p & 3 = 2 (L&1) + e
  {
    const int j = 2 * (lane & 1) + e;
    const I2 w1 = T.tw4[0][j], w2 = T.tw4[1][j], w3 = T.tw4[2][j];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int th = 0; th < 4; th++)
      bfly_mid_raw<SAT>(re[4 * th], im[4 * th], re[4 * th + 1], im[4 * th + 1], re[4 * th + 2], im[4 * th + 2],
                        re[4 * th + 3], im[4 * th + 3], w1, w2, w3);
  }
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int t = 0; t < 16; t++) x[base + 2 * t] = (t & 3) ? pack_hi(re[t], im[t]) : pack_c(re[t], im[t]);
}

ED_HD int rev2(int i) { return ((i & 1) << 1) | ((i >> 1) & 1); }
ED_HD int rev5(int i) {
  return ((i & 1) << 4) | ((i & 2) << 2) | (i & 4) | ((i >> 2) & 2) | ((i >> 4) & 1);
}

// ---- phase C: stage 5 (outputs 0 and 2 of each butterfly), cmpl_mag_q15 --------------------------
// Butterfly j sits at positions 4j..4j+3; its outputs 4j and 4j+2 are bins r = rev8(j) and r + 256.
// Outputs 4j+1, 4j+3 are bins >= 512, which nothing reads (the mel filters end at bin 486;
// audioprocessing.c:157-170).  Trip s (0, 1) of lane L handles the four butterflies
// j = 64 rev2(v) + 2 rev5(L) + s, v = 0..3, i.e. the consecutive bins 4L + 128 s + v (and + 256),
// whose magnitudes are packed four to a word into the two byte planes of this frame's row.
template <bool SAT>
ED_HD void phase_c(int lane, int s, const MfccTables &T, const uint32_t *x, uint32_t *row) {
  const int rho = rev5(lane);
  int m0[4], m2[4];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int v = 0; v < 4; v++) {
    // positions 4j+k: (k = 0, 2) are adjacent words of the even half, (1, 3) of the odd half
    const int a = 34 * (4 * rev2(v) + (rho >> 3)) + 4 * (rho & 7) + 2 * s;
    const I2 ac = *reinterpret_cast<const I2 *>(x + a), bd = *reinterpret_cast<const I2 *>(x + kXchHalf + a);
    const int ar = c_re((uint32_t)ac.x), ai = c_im((uint32_t)ac.x), cr = c_re((uint32_t)ac.y), ci = c_im((uint32_t)ac.y);
    const int br = c_re((uint32_t)bd.x), bi = c_im((uint32_t)bd.x), dr = c_re((uint32_t)bd.y), di = c_im((uint32_t)bd.y);
    const int Rr = qadd16<SAT>(ar, cr), Ri = qadd16<SAT>(ai, ci);
    const int Sr = qsub16<SAT>(ar, cr), Si = qsub16<SAT>(ai, ci);
    const int Tr = qadd16<SAT>(br, dr), Ti = qadd16<SAT>(bi, di);
    const int Ur = qsub16<SAT>(br, dr), Ui = qsub16<SAT>(bi, di);
    mag_q15_x2((Rr + Tr) >> 1, (Ri + Ti) >> 1, (Sr + Ui) >> 1, (Si - Ur) >> 1, T.sqrt_exc, m0[v], m2[v]);
  }
  const int w = lane + 32 * s;
  row[w] = pack4_plane<0>(m0[0], m0[1], m0[2], m0[3]);
  row[w + 64] = pack4_plane<0>(m2[0], m2[1], m2[2], m2[3]);
  row[kSpectPlaneWords + w] = pack4_plane<1>(m0[0], m0[1], m0[2], m0[3]);
  row[kSpectPlaneWords + w + 64] = pack4_plane<1>(m2[0], m2[1], m2[2], m2[3]);
}

// ---- mel filterbank (audioprocessing.c:157-170) as a u8 x u8 tensor-core contraction ---------------
//   D[filter][frame] = sum_bin W[filter][bin] * S[bin][frame],  S = low-byte plane, then high-byte plane
//   mel = (q15)( (D_lo + 256 D_hi) / 128 )        (operands >= 0, so '/' is a shift; the cast wraps)
// m16n8k32 fragments (g = lane >> 2, t = lane & 3): A = T.mel_a[pair][lane]; B word b0 = bins 32ks+4t..+3,
// b1 = bins 32ks+16+4t..+3 of frame g & 3 (columns 4..7 of the tile are never read back).
ED_HD int mel_b_word(int lane, int ks, int plane, int half) {
  return plane * kSpectPlaneWords + ((lane >> 2) & 3) * kSpectRowWords + 8 * ks + 4 * half + (lane & 3);
}
// accumulator element j (0..3) of tile mt -> filter, frame-in-half-group (valid when the frame is < 4)
ED_HD int mel_c_filter(int lane, int mt, int j) { return 16 * mt + (lane >> 2) + 8 * (j >> 1); }
ED_HD int mel_c_frame(int lane, int j) { return 2 * (lane & 3) + (j & 1); }
ED_HD int mel_finish(int lo, int hi) { return (int)(int16_t)(((unsigned)lo + ((unsigned)hi << 8)) >> 7); }

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
