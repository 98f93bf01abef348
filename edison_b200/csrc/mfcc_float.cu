// mfcc_float.cu -- K3: the reference's host-side float MFCCs on the GPU (include/edison_b200.h section 2b).
//
// Reference (audio/edison/mfcc/mfcc_utils.py): mfcc() :134-199 = |FFT| of the raw int16 chunk, first 512 bins, 32 mel bins
// on linspace(0, 8000, 512) -> log(. + 1e-6) -> scipy DCT-II / sqrt(64);  mfcc_mcu() :255-323 = fft/1024, |.|/sqrt(2),
// 513 bins, weights 128 W then /128, optional log, DCT-II / 64.  The reference computes in float64 (numpy).
//
// Two instantiations of one kernel template:
//   R = double  (edison_mfcc_f64_*): the whole chain in fp64 like numpy -- B200 issues DFMA at half the FFMA rate, so this
//               is the default parity path; agrees with numpy to ~1e-9 on ordinary input and stays inside the stated
//               tolerance (max-abs 1e-3, rmse 1e-4) on degenerate frames, where log(mel + 1e-6) sees rounding noise only.
//   R = float   (edison_mfcc_f32_*): fp32 throughout, ~2x less shared-memory traffic; its error is relative to the
//               frame's total energy, so the tolerance holds for frames whose mel-band energies lie within 1e4 of
//               each other (any recorded audio) but not for synthetic pure tones.
//
// One warp per 1024-sample frame.  The real FFT is a 512-point complex FFT of z[n] = x[2n] + j x[2n+1] (the int16 pairs
// are loaded as 32-bit words, coalesced) followed by the real-input split.  512 = 8 x 8 x 8: three radix-8 DIF passes, two
// butterflies per lane per pass, all 16 complex values of a lane in registers, three exchanges through a per-warp
// shared-memory buffer whose layouts are conflict-free (tools/proto/f32_fft_flow.py checks the index flow and the
// wavefront counts).  Magnitudes go to shared memory; the mel filterbank runs 4 lanes per filter, 8 filters at a time;
// log and the 32x32 DCT matrix finish in registers; the 32 coefficients of a frame leave as one coalesced store.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "edison_internal.h"
#include "../../include/edison_b200.h"

namespace edison {

constexpr int kXchStride = 34;                 // complex units per radix-16 output row: 32 + 2 keeps the transposing read conflict-free
constexpr int kXchLen = 16 * kXchStride;       // 544 >= 512
constexpr int kMagLen = 520;
constexpr int kMelWMax = 2304;                 // non-zero weights + chunk padding + the bank alignment of fill_tables

template <typename R> struct alignas(2 * sizeof(R)) Cx { R x, y; };   // one 64- / 128-bit shared-memory access
template <typename R> struct RealTraits;
template <> struct RealTraits<float> { static constexpr int kWarps = 20; };
template <> struct RealTraits<double> { static constexpr int kWarps = 12; };

template <typename R>
struct alignas(16) FloatTables {
  Cx<R> twa[4][16];            // W256^{b p}, p = 1, 2, 4, 8; b = lane >> 1   (after pass A)
  Cx<R> tws[32];               // W1024^{l}                                  (last radix-2 stage + real split)
  R dct_t[32][32];             // [n][k] = scale * 2 cos(pi k (2n+1) / 64)
  R mel_w[kMelWMax];           // non-zero weights filter by filter, times the spectrum scale of the mode (and the 1/2
                               // the split leaves out)
  R log_eps;
  float window[1024];          // mfcc.c mode (fp32 only): Hann window of frame_len samples (mfcc.c:67-68), zero beyond
  R preemph, q_scale;          // mfcc.c mode: pre-emphasis factor (:180-184) and 2^dec_bits (:241)
  int mel_start[32], mel_count[32], mel_off[32];
  int use_log;
  int c_mode;                  // 1: firmware/src/audio/mfcc.c semantics (pre-emphasis, window, FLT_MIN guard, q7 output)
  int frame_len;               // samples read per frame (<= 1024); the rest of the 1024-point transform is zero padding
  int bin_step;                // 1024 / padded frame length: bin i of the reference's FFT is bin i * bin_step here
  int q7_count;                // mfcc.c mode: number of coefficients written (num - offset)
  int mel_count_max;           // host modes: largest (padded) filter length
  int pad_[2];
};

template <typename R>
struct FloatLaunch {
  const int16_t *pcm;
  long long n_clips;
  int frames_per_clip;
  long long clip_stride;
  R *mfcc;                     // [n_clips * frames_per_clip][32]
  int8_t *q7;                  // mfcc.c mode: [n_clips * frames_per_clip][q7_count]
  const FloatTables<R> *tables;
};

template <typename R>
struct alignas(16) WarpFloat {
  Cx<R> xch[kXchLen];
  R mag[kMagLen];
  R mel[32];
  R lm[32];
};
template <typename R>
struct CtaFloat {
  FloatTables<R> T;
  WarpFloat<R> w[RealTraits<R>::kWarps];
};

template <typename R> __device__ __forceinline__ Cx<R> cadd(Cx<R> a, Cx<R> b) { return Cx<R>{a.x + b.x, a.y + b.y}; }
template <typename R> __device__ __forceinline__ Cx<R> csub(Cx<R> a, Cx<R> b) { return Cx<R>{a.x - b.x, a.y - b.y}; }
template <typename R> __device__ __forceinline__ Cx<R> cmul(Cx<R> a, Cx<R> w) {
  return Cx<R>{a.x * w.x - a.y * w.y, a.x * w.y + a.y * w.x};
}
template <typename R> __device__ __forceinline__ Cx<R> mul_mj(Cx<R> a) { return Cx<R>{a.y, -a.x}; }   // a * (-j)

// |z| from the power p = re^2 + im^2
__device__ __forceinline__ float mag_from_power(float p) {   // MUFU.SQRT, relative error <= 2^-22
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(p));
  return r;
}
__device__ __forceinline__ double mag_from_power(double p) {
  // fp32 reciprocal-sqrt seed (relative error ~2^-22), one Newton step in fp64 -> relative error ~1e-13
  if (p < 1e-30) return 0.0;      // |X|^2 of int16 data is 0 or >= ~1e-20 (rounding noise); keeps the seed finite
  float r0;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"((float)p));
  const double r = (double)r0;
  const double s0 = p * r;
  const double e = fma(-s0, s0, p);
  return fma(e, 0.5 * r, s0);
}
__device__ __forceinline__ void load4(const float *p, float (&o)[4]) {
  const float4 v = *reinterpret_cast<const float4 *>(p);
  o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
}
__device__ __forceinline__ void load4(const double *p, double (&o)[4]) {
  const double2 a = *reinterpret_cast<const double2 *>(p), b = *reinterpret_cast<const double2 *>(p + 2);
  o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
}
// the two halves of a 4-element chunk in the order (sw, 2 - sw), sw = 0 or 2; element order is the same for every array
// loaded this way, which is all a dot product needs
__device__ __forceinline__ void load4_swapped(const double *p, int sw, double (&o)[4]) {
  const double2 a = *reinterpret_cast<const double2 *>(p + sw), b = *reinterpret_cast<const double2 *>(p + 2 - sw);
  o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
}
__device__ __forceinline__ void load4_swapped(const float *p, int, float (&o)[4]) { load4(p, o); }
__device__ __forceinline__ float log_r(float x) { return logf(x); }
__device__ __forceinline__ double log_r(double x) { return log(x); }

// W32^j = exp(-2 pi i j / 32), j = 0..7
__device__ constexpr double kW32[8][2] = {
    {1.0, -0.0},
    {0.98078528040323044913, -0.19509032201612826785},
    {0.92387953251128675613, -0.38268343236508977173},
    {0.83146961230254523708, -0.55557023301960222474},
    {0.70710678118654752440, -0.70710678118654752440},
    {0.55557023301960222474, -0.83146961230254523708},
    {0.38268343236508977173, -0.92387953251128675613},
    {0.19509032201612826785, -0.98078528040323044913}};

// 8-point DFT, decimation in frequency, natural order out: v[r] = sum_n v[n] W8^{n r}
template <typename R>
__device__ __forceinline__ void radix8(Cx<R> (&v)[8]) {
  const R s = (R)0.70710678118654752440;
  Cx<R> a[4], b[4];
#pragma unroll
  for (int i = 0; i < 4; i++) { a[i] = cadd(v[i], v[i + 4]); b[i] = csub(v[i], v[i + 4]); }
  b[1] = Cx<R>{(b[1].x + b[1].y) * s, (b[1].y - b[1].x) * s};       // * W8^1
  b[2] = mul_mj(b[2]);                                              // * W8^2
  b[3] = Cx<R>{(b[3].y - b[3].x) * s, -(b[3].x + b[3].y) * s};      // * W8^3
  {
    const Cx<R> e0 = cadd(a[0], a[2]), e1 = csub(a[0], a[2]), e2 = cadd(a[1], a[3]), e3 = mul_mj(csub(a[1], a[3]));
    v[0] = cadd(e0, e2); v[4] = csub(e0, e2); v[2] = cadd(e1, e3); v[6] = csub(e1, e3);
  }
  {
    const Cx<R> e0 = cadd(b[0], b[2]), e1 = csub(b[0], b[2]), e2 = cadd(b[1], b[3]), e3 = mul_mj(csub(b[1], b[3]));
    v[1] = cadd(e0, e2); v[5] = csub(e0, e2); v[3] = cadd(e1, e3); v[7] = csub(e1, e3);
  }
}

// 16-point DFT, decimation in frequency into two radix-8 halves, natural order out: v[r] = sum_n v[n] W16^{n r}
template <typename R>
__device__ __forceinline__ void radix16(Cx<R> (&v)[16]) {
  const R s = (R)0.70710678118654752440, c1 = (R)0.92387953251128675613, s1 = (R)0.38268343236508977173;
  Cx<R> a[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; i++) { a[i] = cadd(v[i], v[i + 8]); b[i] = csub(v[i], v[i + 8]); }
  b[1] = cmul(b[1], Cx<R>{c1, -s1});                                // * W16^1
  b[2] = Cx<R>{(b[2].x + b[2].y) * s, (b[2].y - b[2].x) * s};       // * W16^2 = W8^1
  b[3] = cmul(b[3], Cx<R>{s1, -c1});                                // * W16^3
  b[4] = mul_mj(b[4]);                                              // * W16^4 = -j
  b[5] = cmul(b[5], Cx<R>{-s1, -c1});                               // * W16^5
  b[6] = Cx<R>{(b[6].y - b[6].x) * s, -(b[6].x + b[6].y) * s};      // * W16^6 = W8^3
  b[7] = cmul(b[7], Cx<R>{-c1, -s1});                               // * W16^7
  radix8(a);
  radix8(b);
#pragma unroll
  for (int r = 0; r < 8; r++) { v[2 * r] = a[r]; v[2 * r + 1] = b[r]; }
}

// two adjacent complex values: one 128-bit shared-memory access for float
__device__ __forceinline__ void load_pair(const Cx<float> *p, Cx<float> &a, Cx<float> &b) {
  const float4 v = *reinterpret_cast<const float4 *>(p);
  a = Cx<float>{v.x, v.y}; b = Cx<float>{v.z, v.w};
}
__device__ __forceinline__ void load_pair(const Cx<double> *p, Cx<double> &a, Cx<double> &b) { a = p[0]; b = p[1]; }

// (lo, hi) int16 pair -> two reals without the conversion pipe.
//   float : 0x4B000000 | (x ^ 0x8000)  is 2^23 + x + 2^15
//   double: 0x43300000:(x ^ 0x8000)    is 2^52 + x + 2^15
__device__ __forceinline__ Cx<float> pcm_pair(uint32_t w, float) {
  const uint32_t t = w ^ 0x80008000u;
  const float lo = __uint_as_float(__byte_perm(t, 0x4B00u, 0x5410)) - 8421376.0f;
  const float hi = __uint_as_float(__byte_perm(t, 0x4B00u, 0x5432)) - 8421376.0f;
  return Cx<float>{lo, hi};
}
__device__ __forceinline__ Cx<double> pcm_pair(uint32_t w, double) {
  const uint32_t t = w ^ 0x80008000u;
  const double bias = 4503599627370496.0 + 32768.0;
  const double lo = __hiloint2double(0x43300000, (int)(t & 0xFFFFu)) - bias;
  const double hi = __hiloint2double(0x43300000, (int)(t >> 16)) - bias;
  return Cx<double>{lo, hi};
}

template <bool ALIGNED>
__device__ __forceinline__ void load_frame_words(const int16_t *base, int lane, int frame_len, uint32_t (&w)[16]) {
  if (ALIGNED) {   // base 4-byte aligned and frame_len even
    const uint32_t *p = reinterpret_cast<const uint32_t *>(base);
#pragma unroll
    for (int k = 0; k < 16; k++) w[k] = 2 * (lane + 32 * k) < frame_len ? __ldg(p + lane + 32 * k) : 0u;
  } else {
    const uint16_t *p = reinterpret_cast<const uint16_t *>(base);
#pragma unroll
    for (int k = 0; k < 16; k++) {
      const int n = 2 * (lane + 32 * k);
      const uint32_t lo = n < frame_len ? (uint32_t)__ldg(p + n) : 0u;
      const uint32_t hi = n + 1 < frame_len ? (uint32_t)__ldg(p + n + 1) : 0u;
      w[k] = lo | (hi << 16);
    }
  }
}

template <typename R, bool ALIGNED, bool CMODE>
__global__ void __launch_bounds__(RealTraits<R>::kWarps * 32, 1) mfcc_float_kernel(FloatLaunch<R> P) {
  constexpr int kWarps = RealTraits<R>::kWarps;
  using C = Cx<R>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  CtaFloat<R> &S = *reinterpret_cast<CtaFloat<R> *>(smem_raw);
  {
    const uint32_t *src = reinterpret_cast<const uint32_t *>(P.tables);
    uint32_t *dst = reinterpret_cast<uint32_t *>(&S.T);
    for (int i = threadIdx.x; i < (int)(sizeof(FloatTables<R>) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const FloatTables<R> &T = S.T;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  WarpFloat<R> &W = S.w[warp];
  const long long n_frames = P.n_clips * P.frames_per_clip;
  const long long stride = (long long)gridDim.x * kWarps;
  long long f = (long long)blockIdx.x * kWarps + warp;
  if (f >= n_frames) return;

  auto frame_base = [&](long long fr) {
    const long long clip = fr / P.frames_per_clip;
    return P.pcm + clip * P.clip_stride + (fr - clip * P.frames_per_clip) * (long long)EDISON_FRAME_LEN;
  };
  if (lane < kMagLen - 512) W.mag[512 + lane] = (R)0;    // the padded mel chunks may read past bin 512
  uint32_t pw[16];
  const int frame_len = CMODE ? T.frame_len : 1024;
  load_frame_words<ALIGNED>(frame_base(f), lane, frame_len, pw);

  for (; f < n_frames; f += stride) {
    C v[16];
#pragma unroll
    for (int k = 0; k < 16; k++) v[k] = pcm_pair(pw[k], R());
    if (CMODE) {
      // mfcc.c:176-192: frame[0] = x[0] (unscaled, as the reference has it), frame[i] = (x[i] - 0.97 x[i-1]) / 2^15, then the
      // window; samples beyond frame_len are the zero padding.  x[2n-1] is the neighbouring lane's high half.
      const uint16_t *ps = reinterpret_cast<const uint16_t *>(frame_base(f));
#pragma unroll
      for (int k = 0; k < 16; k++) {
        const int n = lane + 32 * k, s0 = 2 * n;
        if (s0 < frame_len) {
          const R prev = n ? (R)(int16_t)__ldg(ps + s0 - 1) : (R)0;
          const R x0 = v[k].x, x1 = v[k].y;
          const R f0 = n ? (x0 - prev * T.preemph) * (R)(1.0 / 32768.0) : x0;
          const R f1 = (x1 - x0 * T.preemph) * (R)(1.0 / 32768.0);
          v[k].x = f0 * T.window[s0];
          v[k].y = s0 + 1 < frame_len ? f1 * T.window[s0 + 1] : (R)0;
        } else {
          v[k].x = v[k].y = (R)0;
        }
      }
    }
    if (f + stride < n_frames) load_frame_words<ALIGNED>(frame_base(f + stride), lane, frame_len, pw);   // lands during the FFT

    // The 512-point complex FFT of z[n] = x[2n] + j x[2n+1] as 16 x 16 x 2 (n = 32 a + 2 b + c, k = k0 + 16 k1 + 256 k2): two radix-16
    // passes in registers with ONE shared-memory exchange between them and one after; the last radix-2 stage is folded into the
    // real-input split, which reads pairs anyway (tools/proto/fft_16x16x2_flow.py checks this index flow against numpy).  The kernel
    // is bound by shared-memory wavefronts: 8 x 8 x 8 needed three exchanges (228 wavefronts per frame for FFT + split; now 154).
    // ---- pass A: over a (stride 32); lane m = 2 b + c holds z[32 a + m] in v[a]
    radix16(v);
    {
      // twiddle W256^{b k0}, b = lane >> 1: W^1, W^2, W^4, W^8 are read (lane pairs share an entry), the other eleven are products of those
      const int bq = lane >> 1;
      C tw[16];
      tw[1] = T.twa[0][bq]; tw[2] = T.twa[1][bq]; tw[4] = T.twa[2][bq]; tw[8] = T.twa[3][bq];
      tw[3] = cmul(tw[1], tw[2]); tw[5] = cmul(tw[4], tw[1]); tw[6] = cmul(tw[4], tw[2]); tw[7] = cmul(tw[4], tw[3]);
      W.xch[lane] = v[0];
#pragma unroll
      for (int k0 = 1; k0 < 8; k0++) W.xch[kXchStride * k0 + lane] = cmul(v[k0], tw[k0]);
      W.xch[kXchStride * 8 + lane] = cmul(v[8], tw[8]);
#pragma unroll
      for (int k0 = 9; k0 < 16; k0++) W.xch[kXchStride * k0 + lane] = cmul(v[k0], cmul(tw[8], tw[k0 - 8]));
    }
    __syncwarp();
    // ---- pass B: over b; this lane takes k0 = lane >> 1 and c = lane & 1 (row stride 34: the 16 lanes of a wavefront hit 16 bank pairs)
    {
      const C *row = W.xch + kXchStride * (lane >> 1) + (lane & 1);
#pragma unroll
      for (int b = 0; b < 16; b++) v[b] = row[2 * b];
    }
    __syncwarp();
    radix16(v);
    // Y2[k0, k1; c] at element 2 (k0 + 16 k1) + c = lane + 32 k1: the pair (c = 0, c = 1) of one k is one 128-bit read below
#pragma unroll
    for (int k1 = 0; k1 < 16; k1++) W.xch[lane + 32 * k1] = v[k1];
    __syncwarp();
    // ---- last stage + real-input split.  For k = lane + 32 j < 128: (P, Q) = Y2[k; 0..1], Z[k] = P + W512^k Q, Z[k + 256] = P - W512^k Q;
    //      the same for k' = 256 - k gives Z[256 - k] and Z[512 - k]; then the split on the pairs (k, 512 - k) and (256 - k, 256 + k).
    //      Magnitudes are 2|X| (the 1/2 lives in mel_w).  W1024^k = W1024^lane * W32^j, W512^k is its square, W512^{256-k} = -conj(W512^k),
    //      W1024^{256-k} = -j conj(W1024^k): one table read per frame.
    auto split_pair = [&](C a, C bc, C w, int ka, int kb) {      // a = Z[ka], bc = Z[kb] with kb = 512 - ka, w = W1024^ka
      const R er = a.x + bc.x, ei = a.y - bc.y;        // A + conj(B)
      const R dr = a.x - bc.x, di = a.y + bc.y;        // A - conj(B)
      const R tr = w.x * di + w.y * dr, ti = w.y * di - w.x * dr;   // W^k * (-j D)
      const R x1r = er + tr, x1i = ei + ti, x2r = er - tr, x2i = ei - ti;
      W.mag[ka] = mag_from_power(x1r * x1r + x1i * x1i);
      W.mag[kb] = mag_from_power(x2r * x2r + x2i * x2i);
    };
    const C ws0 = T.tws[lane];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int k = lane + 32 * j, kp = 256 - k;       // lane 0, j = 0 computes rubbish here (kp = 256) and repairs it below
      const C w = j ? cmul(ws0, C{(R)kW32[j][0], (R)kW32[j][1]}) : ws0;
      const C w2 = cmul(w, w);
      C p, q, pp, qp;
      load_pair(W.xch + 2 * k, p, q);
      load_pair(W.xch + 2 * kp, pp, qp);
      const C t = cmul(w2, q), tp = cmul(C{-w2.x, w2.y}, qp);
      const C zk = cadd(p, t), zk256 = csub(p, t), zkp = cadd(pp, tp), z512k = csub(pp, tp);
      split_pair(zk, z512k, w, k, 512 - k);
      split_pair(zkp, zk256, C{-w.y, -w.x}, kp, 256 + k);
    }
    if (lane == 0) {
      C p, q;
      load_pair(W.xch, p, q);                          // k = 0: Z[0] = P + Q pairs with itself (Z[512] = Z[0]); Z[256] = P - Q
      const C z0 = cadd(p, q), z256 = csub(p, q);
      split_pair(z0, z0, C{(R)1, (R)0}, 0, 512);
      W.mag[256] = (R)2 * mag_from_power(z256.x * z256.x + z256.y * z256.y);
      load_pair(W.xch + 256, p, q);                    // k = 128: W512^128 = -j, W1024^128 = exp(-j pi / 4)
      const C t = mul_mj(q);
      split_pair(cadd(p, t), csub(p, t), C{(R)0.70710678118654752440, (R)-0.70710678118654752440}, 128, 384);
    }
    __syncwarp();
    // ---- mel: 4 lanes per filter, 8 filters per round.  Each lane takes aligned chunks of 4 bins (128-bit reads of the
    //      magnitudes and of the zero-padded weights); the mfcc.c mode strides through the bins and stays scalar.
    //      (8 lanes per filter would make every quarter-warp read 128 contiguous bytes -- no bank conflicts, 92 instead of 143
    //      wavefronts per frame -- but doubles the rounds, each a dependent load-FMA-shuffle chain: measured 22 % slower.)
    if (CMODE) {
      const int q = lane & 3;
#pragma unroll
      for (int round = 0; round < 4; round++) {
        const int m = 8 * round + (lane >> 2);
        const int cnt = T.mel_count[m];
        const R *wt = T.mel_w + T.mel_off[m];
        const int bs = T.bin_step;
        const R *mg = W.mag + T.mel_start[m] * bs;
        R acc = (R)0;
        for (int i = q; i < cnt; i += 4) acc = fma(mg[i * bs], wt[i], acc);
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        if (q == 0) W.mel[m] = acc;
      }
    } else {
      // the four rounds advance together, so that eight independent 128-bit loads are in flight per step (one round at a
      // time is a chain of dependent load-FMA steps, and this part of the kernel waits on shared-memory latency)
      const int q = lane & 3;
      int cnt[4];
      const R *wt[4], *mg[4];
      R acc[4];
#pragma unroll
      for (int r = 0; r < 4; r++) {
        const int m = 8 * r + (lane >> 2);
        cnt[r] = T.mel_count[m];
        wt[r] = T.mel_w + T.mel_off[m];
        mg[r] = W.mag + T.mel_start[m];
        acc[r] = (R)0;
      }
      // Bank conflicts: a quarter-warp (one wavefront of a 128-bit load) holds two filters with unrelated start bins.
      // double: a 4-bin chunk is two 16-byte loads, and the odd filter of the pair issues them in the other order, so the
      // eight lanes always read eight different 16-byte bank groups (fp64 +8 %).  float: a chunk is one load; taking every
      // second chunk with opposite parities in the two filters is conflict-free too, but splits short filters over two
      // steps and measured 8 % slower -- the float kernel keeps the contiguous order and its conflicts.
      const int cmax = T.mel_count_max;
      const int sw = 2 * ((lane >> 2) & 1);
      for (int i = 4 * q; i < cmax; i += 16) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
          if (i < cnt[r]) {
            R a[4], w[4];
            load4_swapped(mg[r] + i, sw, a);
            load4_swapped(wt[r] + i, sw, w);
            acc[r] = fma(a[0], w[0], acc[r]); acc[r] = fma(a[1], w[1], acc[r]);
            acc[r] = fma(a[2], w[2], acc[r]); acc[r] = fma(a[3], w[3], acc[r]);
          }
        }
      }
#pragma unroll
      for (int r = 0; r < 4; r++) {
        acc[r] += __shfl_xor_sync(0xffffffffu, acc[r], 1);
        acc[r] += __shfl_xor_sync(0xffffffffu, acc[r], 2);
        if (q == 0) W.mel[8 * r + (lane >> 2)] = acc[r];
      }
    }
    __syncwarp();
    {
      R x = W.mel[lane];
      if (CMODE) x = log_r(x == (R)0 ? (R)1.17549435e-38 : x);     // FLT_MIN for empty bands, mfcc.c:225-231
      else if (T.use_log) x = log_r(x + T.log_eps);
      W.lm[lane] = x;
    }
    __syncwarp();
    {
      R a4[4] = {(R)0, (R)0, (R)0, (R)0};   // four chains instead of one of 32 dependent FMAs
      if (CMODE) {
#pragma unroll
        for (int n = 0; n < 32; n += 4) {
          R l[4];
          load4(W.lm + n, l);
#pragma unroll
          for (int u = 0; u < 4; u++) a4[u] = fma(l[u], T.dct_t[n + u][lane], a4[u]);
        }
      } else {
        // DCT-II over 32 points: row 31 - n of the matrix is (-1)^k times row n, so half of the table is read and the
        // inputs are folded first, lm[n] + lm[31 - n] for even k (= lane), lm[n] - lm[31 - n] for odd k
        const R sgn = (lane & 1) ? (R)-1 : (R)1;
#pragma unroll
        for (int n = 0; n < 16; n += 4) {
          R l[4], h[4];
          load4(W.lm + n, l);
          load4(W.lm + 28 - n, h);
#pragma unroll
          for (int u = 0; u < 4; u++) a4[u] = fma(fma(sgn, h[3 - u], l[u]), T.dct_t[n + u][lane], a4[u]);
        }
      }
      const R acc = (a4[0] + a4[1]) + (a4[2] + a4[3]);
      if (CMODE) {
        // mfcc.c:240-250: sum * 2^dec_bits, round() (half away from zero), saturate to q7
        if (lane < T.q7_count) {
          R r = acc * T.q_scale;
          r = r < (R)0 ? -floor((R)0.5 - r) : floor(r + (R)0.5);
          P.q7[f * T.q7_count + lane] = (int8_t)(r >= (R)127 ? 127 : (r <= (R)-128 ? -128 : (int)r));
        }
      } else {
        P.mfcc[f * 32 + lane] = acc;
      }
    }
    __syncwarp();
  }
}

// ---- tables ------------------------------------------------------------------------------------------------------------
// gen_mel_weight_matrix (mfcc_utils.py:36-73) with the constants of audio/config.py:16-21, 33-34; everything is built in
// double and rounded once to R
template <typename R>
static int fill_tables(FloatTables<R> &T, int mode) {
  memset(&T, 0, sizeof T);
  const double tau = 6.283185307179586476925286766559;
  for (int pi = 0; pi < 4; pi++)
    for (int b = 0; b < 16; b++) {
      const double a = -tau * (double)((b << pi) % 256) / 256.0;
      T.twa[pi][b] = Cx<R>{(R)cos(a), (R)sin(a)};
    }
  for (int l = 0; l < 32; l++) {
    const double a = -tau * (double)l / 1024.0;
    T.tws[l] = Cx<R>{(R)cos(a), (R)sin(a)};
  }
  const int n_bins = mode == EDISON_MFCC_F32_HOST ? 512 : 513;
  const double spec_scale = mode == EDISON_MFCC_F32_HOST ? 1.0 : 1.0 / 1024.0 / sqrt(2.0);
  const double dct_scale = mode == EDISON_MFCC_F32_HOST ? 1.0 / sqrt(64.0) : 1.0 / 64.0;
  T.use_log = mode != EDISON_MFCC_F32_MCU;
  T.log_eps = (R)1e-6;
  T.frame_len = 1024;
  T.bin_step = 1;
  auto mel = [](double f) { return 1127.0 * log(1.0 + f / 700.0); };
  const double e_lo = mel(80.0), e_hi = mel(7600.0);
  double edges[34];
  for (int i = 0; i < 34; i++) edges[i] = e_lo + (e_hi - e_lo) * (double)i / 33.0;
  int off = 0;
  for (int m = 0; m < 32; m++) {
    int start = -1, count = 0;
    std::vector<double> w;
    for (int b = 1; b < n_bins; b++) {
      const double x = mel(8000.0 * (double)b / (double)(n_bins - 1));
      const double lower = (x - edges[m]) / (edges[m + 1] - edges[m]);
      const double upper = (edges[m + 2] - x) / (edges[m + 2] - edges[m + 1]);
      const double v = fmax(0.0, fmin(lower, upper));
      if (v > 0.0) {
        if (start < 0) start = b;
        if (b != start + count) return -1;   // a filter's support is one run of bins
        w.push_back(v);
        count++;
      }
    }
    if (start < 0) start = 0;
    const int a_start = start & ~3;                                  // chunks of 4 bins, 16-byte aligned in mag[]
    const int a_count = ((start + count - a_start) + 3) & ~3;
    if (sizeof(R) == 4) {
      // fp32: the two filters that share a quarter-warp in the mel loop (m even, m odd: lanes 8p..8p+3 and 8p+4..8p+7 of one 128-bit
      // load wavefront) each read 64 contiguous bytes of weights per step; starting them 64 bytes apart modulo 128 makes those reads
      // conflict-free whatever the filters' lengths (the magnitudes' alignment is fixed by the filterbank and keeps its conflicts)
      const int want = (m & 1) ? 16 : 0;
      off += ((want - off) % 32 + 32) % 32;
    }
    if (off + a_count > kMelWMax || a_start + a_count > kMagLen) return -1;
    T.mel_start[m] = a_start; T.mel_count[m] = a_count; T.mel_off[m] = off;
    if (a_count > T.mel_count_max) T.mel_count_max = a_count;
    for (int i = 0; i < count; i++) T.mel_w[off + (start - a_start) + i] = (R)(w[i] * spec_scale * 0.5);
    off += a_count;
  }
  for (int n = 0; n < 32; n++)
    for (int k = 0; k < 32; k++)
      T.dct_t[n][k] = (R)(dct_scale * 2.0 * cos(3.14159265358979323846 * (double)k * (double)(2 * n + 1) / 64.0));
  return 0;
}

template <typename R>
struct FloatState {
  FloatTables<R> *tables[3] = {nullptr, nullptr, nullptr};   // device, built on first use (under the ctx lock)
  bool configured = false;
};
static FloatState<float> g_f32;
static FloatState<double> g_f64;
template <typename R> static FloatState<R> &state();
template <> FloatState<float> &state<float>() { return g_f32; }
template <> FloatState<double> &state<double>() { return g_f64; }

template <typename R>
static int float_tables(int mode, const FloatTables<R> **out) {
  FloatState<R> &st = state<R>();
  if (mode < 0 || mode > 2) return ctx_fail(EDISON_E_ARG, "float MFCC mode %d unknown", mode);
  if (!st.configured) {
    cudaError_t e = cudaFuncSetAttribute(mfcc_float_kernel<R, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)sizeof(CtaFloat<R>));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(mfcc_float_kernel<R, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(CtaFloat<R>));
    if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "mfcc_float shared memory opt-in: %s", cudaGetErrorString(e));
    st.configured = true;
  }
  if (!st.tables[mode]) {
    FloatTables<R> *h = new FloatTables<R>;
    if (fill_tables(*h, mode)) { delete h; return ctx_fail(EDISON_E_DATA, "mel table construction failed"); }
    FloatTables<R> *d = nullptr;
    cudaError_t e = cudaMalloc(&d, sizeof *h);
    if (e == cudaSuccess) e = cudaMemcpy(d, h, sizeof *h, cudaMemcpyHostToDevice);
    delete h;
    if (e != cudaSuccess) { if (d) cudaFree(d); return ctx_fail(EDISON_E_CUDA, "float MFCC tables: %s", cudaGetErrorString(e)); }
    st.tables[mode] = d;
  }
  *out = st.tables[mode];
  return EDISON_OK;
}

template <typename R>
static int launch_float(const CtxView &c, const int16_t *d_pcm, size_t n_clips, size_t fpc, size_t stride, R *d_out, int mode,
                        cudaStream_t st) {
  constexpr int kWarps = RealTraits<R>::kWarps;
  const FloatTables<R> *tb;
  int rc = float_tables<R>(mode, &tb);
  if (rc) return rc;
  const long long n_frames = (long long)n_clips * (long long)fpc;
  if (n_frames == 0) return EDISON_OK;
  FloatLaunch<R> L{d_pcm, (long long)n_clips, (int)fpc, (long long)stride, d_out, nullptr, tb};
  const long long ctas = (n_frames + kWarps - 1) / kWarps;
  const unsigned grid = (unsigned)(ctas < c.sm_count ? ctas : c.sm_count);
  const bool aligned = ((uintptr_t)d_pcm % 4 == 0) && (stride % 2 == 0);
  if (aligned) mfcc_float_kernel<R, true, false><<<grid, kWarps * 32, sizeof(CtaFloat<R>), st>>>(L);
  else mfcc_float_kernel<R, false, false><<<grid, kWarps * 32, sizeof(CtaFloat<R>), st>>>(L);
  count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "mfcc_float_kernel: %s", cudaGetErrorString(e));
  return EDISON_OK;
}

void mfcc_float_release(void) {
  for (int m = 0; m < 3; m++) {
    if (g_f32.tables[m]) cudaFree(g_f32.tables[m]);
    if (g_f64.tables[m]) cudaFree(g_f64.tables[m]);
    g_f32.tables[m] = nullptr;
    g_f64.tables[m] = nullptr;
  }
}

static bool float_dev_ptr(const void *p) {
  cudaPointerAttributes a;
  if (!p || cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

template <typename R>
static int clips_dev(const int16_t *d_pcm, size_t n_clips, size_t fpc, size_t stride, R *d_mfcc, int mode, void *stream) {
  DeviceScope device_scope_;
  if ((n_clips && fpc) && (!d_pcm || !d_mfcc)) return ctx_fail(EDISON_E_ARG, "float MFCC: null pointer");
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  if (!rc) rc = launch_float<R>(c, d_pcm, n_clips, fpc, stride, d_mfcc, mode, stream ? (cudaStream_t)stream : c.stream);
  ctx_unlock();
  return rc;
}

template <typename R>
static int clips_any(const int16_t *pcm, size_t n_clips, size_t fpc, size_t stride, R *mfcc, int mode) {
  DeviceScope device_scope_;
  const size_t n_frames = n_clips * fpc;
  if (n_frames && (!pcm || !mfcc)) return ctx_fail(EDISON_E_ARG, "float MFCC: null pointer");
  if (float_dev_ptr(pcm)) {
    int rc = clips_dev<R>(pcm, n_clips, fpc, stride, mfcc, mode, nullptr);
    return rc ? rc : edison_sync(nullptr);
  }
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  int16_t *d_pcm = nullptr;
  R *d_out = nullptr;
  if (!rc && n_frames) {
    const size_t samples = (n_clips - 1) * stride + fpc * EDISON_FRAME_LEN;
    cudaError_t e = cudaMalloc(&d_pcm, samples * sizeof(int16_t));
    if (e == cudaSuccess) e = cudaMalloc(&d_out, n_frames * 32 * sizeof(R));
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_pcm, pcm, samples * sizeof(int16_t), cudaMemcpyHostToDevice, c.stream);
    if (e != cudaSuccess) rc = ctx_fail(EDISON_E_NOMEM, "float MFCC staging: %s", cudaGetErrorString(e));
    if (!rc) rc = launch_float<R>(c, d_pcm, n_clips, fpc, stride, d_out, mode, c.stream);
    if (!rc) {
      e = cudaMemcpyAsync(mfcc, d_out, n_frames * 32 * sizeof(R), cudaMemcpyDeviceToHost, c.stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
      if (e != cudaSuccess) rc = ctx_fail(EDISON_E_CUDA, "float MFCC: %s", cudaGetErrorString(e));
    }
    if (d_pcm) cudaFree(d_pcm);
    if (d_out) cudaFree(d_out);
  }
  ctx_unlock();
  return rc;
}

// ---- firmware/src/audio/mfcc.c: table construction in float32 exactly as mfcc_create / create_mel_fbank /
// create_dct_matrix do it (mfcc.c:47-171), so that bin membership of the filters matches the reference build on this host
static int fill_tables_c(FloatTables<float> &T, int num, int offset, int frame_len, int dec_bits, float preemph) {
  if (fill_tables(T, EDISON_MFCC_F32_HOST)) return -1;     // FFT twiddles; everything else is overwritten below
  const int padded = (int)powf(2, ceilf((logf((float)frame_len) / logf(2))));   // mfcc.c:60
  if (frame_len < 2 || padded > 1024 || num < 1 || num > 26 || offset < 0 || offset >= num || dec_bits < 0 || dec_bits > 24) return -2;
  memset(T.dct_t, 0, sizeof T.dct_t);
  memset(T.mel_w, 0, sizeof T.mel_w);
  memset(T.window, 0, sizeof T.window);
  T.c_mode = 1; T.use_log = 1; T.frame_len = frame_len; T.bin_step = 1024 / padded; T.q7_count = num - offset;
  T.preemph = preemph; T.q_scale = (float)(1 << dec_bits);
  for (int i = 0; i < frame_len; i++)
    T.window[i] = 0.5f - 0.5f * cosf((float)6.283185307179586476925286766559005 * ((float)i) / (frame_len));   // :67-68
  auto mel_scale = [](float f) { return 1127.0f * logf(1.0f + f / 700.0f); };                                 // mfcc.h:53-55
  const int nb = padded / 2;
  const float width = ((float)16000) / padded;
  const float mel_lo = mel_scale(20), mel_hi = mel_scale(4000);
  const float delta = (mel_hi - mel_lo) / (26 + 1);
  int off = 0;
  for (int bin = 0; bin < 32; bin++) {
    T.mel_start[bin] = 0; T.mel_count[bin] = 0; T.mel_off[bin] = off;
    if (bin >= 26) continue;
    const float left = mel_lo + bin * delta, center = mel_lo + (bin + 1) * delta, right = mel_lo + (bin + 2) * delta;
    int first = -1, last = -1;
    std::vector<float> w(nb, 0.0f);
    for (int i = 0; i < nb; i++) {
      const float mel = mel_scale(width * i);
      if (mel > left && mel < right) {
        w[i] = mel <= center ? (mel - left) / (center - left) : (right - mel) / (right - center);
        if (first == -1) first = i;
        last = i;
      }
    }
    if (first < 0) continue;          // (the reference would index [-1] here; no configuration <= 1024 reaches it)
    const int count = last - first + 1;
    if (off + count > kMelWMax) return -3;
    T.mel_start[bin] = first; T.mel_count[bin] = count;
    for (int i = 0; i < count; i++) T.mel_w[off + i] = w[first + i] * 0.5f;
    off += count;
  }
  const float normalizer = sqrtf(2.0f / (float)26);                                                           // :101-115
  for (int k = offset; k < num; k++)
    for (int n = 0; n < 26; n++) T.dct_t[n][k - offset] = normalizer * cosf(((float)3.14159265358979323846264338327950288) / 26 * (n + 0.5f) * k);
  return 0;
}

}  // namespace edison

struct edison_mfccf {
  edison::FloatTables<float> *d_tables;
  int n_out, frame_len;
};

namespace edison {

static int mfccf_launch(const CtxView &c, edison_mfccf *h, const int16_t *d_pcm, size_t n_frames, size_t hop, int8_t *d_out,
                        cudaStream_t st) {
  constexpr int kWarps = RealTraits<float>::kWarps;
  static bool configured = false;      // under the ctx lock
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(mfcc_float_kernel<float, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)sizeof(CtaFloat<float>));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(mfcc_float_kernel<float, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)sizeof(CtaFloat<float>));
    if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "mfcc_float (mfcc.c mode) shared memory opt-in: %s", cudaGetErrorString(e));
    configured = true;
  }
  if (!n_frames) return EDISON_OK;
  FloatLaunch<float> L{d_pcm, (long long)n_frames, 1, (long long)hop, nullptr, d_out, h->d_tables};
  const long long ctas = ((long long)n_frames + kWarps - 1) / kWarps;
  const unsigned grid = (unsigned)(ctas < c.sm_count ? ctas : c.sm_count);
  const bool aligned = ((uintptr_t)d_pcm % 4 == 0) && (hop % 2 == 0) && (h->frame_len % 2 == 0);
  if (aligned) mfcc_float_kernel<float, true, true><<<grid, kWarps * 32, sizeof(CtaFloat<float>), st>>>(L);
  else mfcc_float_kernel<float, false, true><<<grid, kWarps * 32, sizeof(CtaFloat<float>), st>>>(L);
  count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "mfcc_float_kernel (mfcc.c mode): %s", cudaGetErrorString(e));
  return EDISON_OK;
}

}  // namespace edison

using namespace edison;

extern "C" {

int edison_mfccf_create(int num_mfcc_features, int feature_offset, int frame_len, int mfcc_dec_bits, float preemph,
                        edison_mfccf **out) {
  DeviceScope device_scope_;
  if (!out) return ctx_fail(EDISON_E_ARG, "edison_mfccf_create: null out");
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  if (!rc) {
    FloatTables<float> *h = new FloatTables<float>;
    const int r = fill_tables_c(*h, num_mfcc_features, feature_offset, frame_len, mfcc_dec_bits, preemph);
    if (r) rc = ctx_fail(EDISON_E_ARG, "edison_mfccf_create(%d, %d, %d, %d): unsupported configuration (frames pad to <= 1024, <= 26 "
                         "coefficients)", num_mfcc_features, feature_offset, frame_len, mfcc_dec_bits);
    FloatTables<float> *d = nullptr;
    if (!rc) {
      cudaError_t e = cudaMalloc(&d, sizeof *h);
      if (e == cudaSuccess) e = cudaMemcpy(d, h, sizeof *h, cudaMemcpyHostToDevice);
      if (e != cudaSuccess) { if (d) cudaFree(d); rc = ctx_fail(EDISON_E_CUDA, "edison_mfccf_create: %s", cudaGetErrorString(e)); }
    }
    delete h;
    if (!rc) *out = new edison_mfccf{d, num_mfcc_features - feature_offset, frame_len};
  }
  ctx_unlock();
  return rc;
}

void edison_mfccf_delete(edison_mfccf *h) {
  DeviceScope device_scope_;
  if (!h) return;
  cudaDeviceSynchronize();
  if (h->d_tables) cudaFree(h->d_tables);
  delete h;
}

int edison_mfccf_n_out(const edison_mfccf *h) { return h ? h->n_out : 0; }

int edison_mfccf_compute_batch_dev(edison_mfccf *h, const int16_t *d_pcm, size_t n_frames, size_t hop_samples, int8_t *d_out,
                                   void *stream) {
  DeviceScope device_scope_;
  if (!h || (n_frames && (!d_pcm || !d_out || !hop_samples))) return ctx_fail(EDISON_E_ARG, "edison_mfccf_compute_batch_dev: bad arguments");
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  if (!rc) rc = mfccf_launch(c, h, d_pcm, n_frames, hop_samples, d_out, stream ? (cudaStream_t)stream : c.stream);
  ctx_unlock();
  return rc;
}

int edison_mfccf_compute_batch(edison_mfccf *h, const int16_t *pcm, size_t n_frames, size_t hop_samples, int8_t *out) {
  DeviceScope device_scope_;
  if (!h || (n_frames && (!pcm || !out || !hop_samples))) return ctx_fail(EDISON_E_ARG, "edison_mfccf_compute_batch: bad arguments");
  if (float_dev_ptr(pcm)) {
    int rc = edison_mfccf_compute_batch_dev(h, pcm, n_frames, hop_samples, out, nullptr);
    return rc ? rc : edison_sync(nullptr);
  }
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  int16_t *d_pcm = nullptr;
  int8_t *d_out = nullptr;
  if (!rc && n_frames) {
    const size_t samples = (n_frames - 1) * hop_samples + (size_t)h->frame_len;
    cudaError_t e = cudaMalloc(&d_pcm, samples * sizeof(int16_t));
    if (e == cudaSuccess) e = cudaMalloc(&d_out, n_frames * (size_t)h->n_out);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_pcm, pcm, samples * sizeof(int16_t), cudaMemcpyHostToDevice, c.stream);
    if (e != cudaSuccess) rc = ctx_fail(EDISON_E_NOMEM, "edison_mfccf_compute_batch staging: %s", cudaGetErrorString(e));
    if (!rc) rc = mfccf_launch(c, h, d_pcm, n_frames, hop_samples, d_out, c.stream);
    if (!rc) {
      e = cudaMemcpyAsync(out, d_out, n_frames * (size_t)h->n_out, cudaMemcpyDeviceToHost, c.stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
      if (e != cudaSuccess) rc = ctx_fail(EDISON_E_CUDA, "edison_mfccf_compute_batch: %s", cudaGetErrorString(e));
    }
    if (d_pcm) cudaFree(d_pcm);
    if (d_out) cudaFree(d_out);
  }
  ctx_unlock();
  return rc;
}

int edison_mfcc_f64_clips_dev(const int16_t *d_pcm, size_t n_clips, size_t frames_per_clip, size_t clip_stride_samples,
                              double *d_mfcc, int mode, void *stream) {
  return clips_dev<double>(d_pcm, n_clips, frames_per_clip, clip_stride_samples, d_mfcc, mode, stream);
}
int edison_mfcc_f64_clips(const int16_t *pcm, size_t n_clips, size_t frames_per_clip, size_t clip_stride_samples, double *mfcc,
                          int mode) {
  return clips_any<double>(pcm, n_clips, frames_per_clip, clip_stride_samples, mfcc, mode);
}
int edison_mfcc_f64_batch(const int16_t *pcm, size_t n_frames, double *mfcc, int mode) {
  return clips_any<double>(pcm, n_frames, 1, EDISON_FRAME_LEN, mfcc, mode);
}
int edison_mfcc_f32_clips_dev(const int16_t *d_pcm, size_t n_clips, size_t frames_per_clip, size_t clip_stride_samples,
                              float *d_mfcc, int mode, void *stream) {
  return clips_dev<float>(d_pcm, n_clips, frames_per_clip, clip_stride_samples, d_mfcc, mode, stream);
}
int edison_mfcc_f32_clips(const int16_t *pcm, size_t n_clips, size_t frames_per_clip, size_t clip_stride_samples, float *mfcc,
                          int mode) {
  return clips_any<float>(pcm, n_clips, frames_per_clip, clip_stride_samples, mfcc, mode);
}
int edison_mfcc_f32_batch(const int16_t *pcm, size_t n_frames, float *mfcc, int mode) {
  return clips_any<float>(pcm, n_frames, 1, EDISON_FRAME_LEN, mfcc, mode);
}

}  // extern "C"
