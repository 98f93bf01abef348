// q15_arith.cuh -- exact Cortex-M4 fixed-point arithmetic, one 16-bit lane per int32.
//
// The reference firmware is built with -DARM_MATH_CM4 -DARM_MATH_DSP (firmware/Makefile:30), so
// CMSIS-DSP takes its packed-SIMD branches.  Every helper here states the M4 instruction (or CMSIS
// function) it reproduces bit for bit; values are kept sign-extended in 32-bit registers, which
// is how a GPU wants them (one IADD3/VIMNMX/IMAD per lane).
//
// The code is plain C++ and compiles for the device (nvcc) and, for the index-math emulation
// harness under tests/host_emu/, for the host (g++).  The product only ever runs the device build.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define ED_HD __host__ __device__ __forceinline__
#else
#define ED_HD inline
#endif

namespace edison {

struct alignas(8) I2 { int x, y; };
struct alignas(16) I4 { int x, y, z, w; };

ED_HD int imin(int a, int b) { return a < b ? a : b; }
ED_HD int imax(int a, int b) { return a > b ? a : b; }

// QADD16 / QSUB16 on one lane (arm_math.h:1368-1697 host emulations; ARMv7-M ARM A7.7.107).
// SAT=false is used where interval analysis proves that saturation cannot trigger.
template <bool SAT> ED_HD int qadd16(int a, int b) {
#if defined(__CUDA_ARCH__)
  if (SAT) return min(__viaddmax_s32(a, b, -32768), 32767);  // VIADDMNMX + VIMNMX
  return a + b;
#else
  int s = a + b;
  if (SAT) s = imin(imax(s, -32768), 32767);
  return s;
#endif
}
template <bool SAT> ED_HD int qsub16(int a, int b) {
#if defined(__CUDA_ARCH__)
  if (SAT) return min(__viaddmax_s32(a, -b, -32768), 32767);
  return a - b;
#else
  int s = a - b;
  if (SAT) s = imin(imax(s, -32768), 32767);
  return s;
#endif
}

// Complex multiply by a q15 twiddle the way the SIMD branch does it
// (arm_cfft_radix4_q15.c:252-256):  re' = SMUAD(C, V) >> 16,  im' = SMUSDX(C, V) >> 16.
// |co*re + si*im| <= 32768 * 65535 < 2^31, so plain int arithmetic never wraps.
// The shift is issued as IMAD.HI (x * 65536 >> 32) on the device: the kernel is bound by the ALU pipe
// (shifts, logic, min/max), the FMA pipe has room, and both give floor(x / 65536).
ED_HD int asr16(int x) {
#if defined(__CUDA_ARCH__)
  return __mulhi(x, 65536);
#else
  return x >> 16;
#endif
}
ED_HD void tw_mul(int co, int si, int re, int im, int &ore, int &oim) {
  ore = asr16(co * re + si * im);
  oim = asr16(co * im - si * re);
}
// the same without the final shift: X = co*re + si*im, Y = co*im - si*re; pack_hi(X, Y) packs
// (X >> 16, Y >> 16) into one word with a single byte permute
ED_HD void tw_mul_raw(int co, int si, int re, int im, int &X, int &Y) {
  X = co * re + si * im;
  Y = co * im - si * re;
}
ED_HD uint32_t pack_hi(int X, int Y) {
#if defined(__CUDA_ARCH__)
  return __byte_perm((uint32_t)X, (uint32_t)Y, 0x7632);
#else
  return (((uint32_t)X >> 16) & 0xffffu) | ((uint32_t)Y & 0xffff0000u);
#endif
}

// byte plane k (0 = low bytes, 1 = high bytes) of four non-negative q15 values, a in the lowest byte: three PRMT
template <int K> ED_HD uint32_t pack4_plane(int a, int b, int c, int d) {
#if defined(__CUDA_ARCH__)
  const uint32_t ab = __byte_perm((uint32_t)a, (uint32_t)b, K ? 0x0051 : 0x0040);
  const uint32_t cd = __byte_perm((uint32_t)c, (uint32_t)d, K ? 0x0051 : 0x0040);
  return __byte_perm(ab, cd, 0x5410);
#else
  const int sh = 8 * K;
  return ((uint32_t)(a >> sh) & 255u) | (((uint32_t)(b >> sh) & 255u) << 8) | (((uint32_t)(c >> sh) & 255u) << 16) |
         (((uint32_t)(d >> sh) & 255u) << 24);
#endif
}

// One radix-4 DIF butterfly of a MIDDLE stage (arm_cfft_radix4_q15.c:333-479).
// a, b, c, d are x[i0], x[i0+n2], x[i0+2n2], x[i0+3n2]; w1/w2/w3 = W^ic, W^2ic, W^3ic.
template <bool SAT>
ED_HD void bfly_mid(int &ar, int &ai, int &br, int &bi, int &cr, int &ci, int &dr, int &di,
                    I2 w1, I2 w2, I2 w3) {
  int Rr = qadd16<SAT>(ar, cr), Ri = qadd16<SAT>(ai, ci);
  int Sr = qsub16<SAT>(ar, cr), Si = qsub16<SAT>(ai, ci);
  int Tr = qadd16<SAT>(br, dr), Ti = qadd16<SAT>(bi, di);
  int Ur = qsub16<SAT>(br, dr), Ui = qsub16<SAT>(bi, di);
  ar = (Rr + Tr) >> 2;  // SHADD16 twice (:372-374): floor of floor == one shift by 2
  ai = (Ri + Ti) >> 2;
  tw_mul(w2.x, w2.y, (Rr - Tr) >> 1, (Ri - Ti) >> 1, br, bi);  // SHSUB16 :378
  tw_mul(w1.x, w1.y, (Sr + Ui) >> 1, (Si - Ur) >> 1, cr, ci);  // SHSAX :415
  tw_mul(w3.x, w3.y, (Sr - Ui) >> 1, (Si + Ur) >> 1, dr, di);  // SHASX :412
}

// bfly_mid whose three twiddled outputs are left unshifted (for the last stage of a phase, where they
// go straight to shared memory through pack_hi)
template <bool SAT>
ED_HD void bfly_mid_raw(int &ar, int &ai, int &bX, int &bY, int &cX, int &cY, int &dX, int &dY, I2 w1, I2 w2, I2 w3) {
  const int br = bX, bi = bY, cr = cX, ci = cY, dr = dX, di = dY;
  int Rr = qadd16<SAT>(ar, cr), Ri = qadd16<SAT>(ai, ci);
  int Sr = qsub16<SAT>(ar, cr), Si = qsub16<SAT>(ai, ci);
  int Tr = qadd16<SAT>(br, dr), Ti = qadd16<SAT>(bi, di);
  int Ur = qsub16<SAT>(br, dr), Ui = qsub16<SAT>(bi, di);
  ar = (Rr + Tr) >> 2;
  ai = (Ri + Ti) >> 2;
  tw_mul_raw(w2.x, w2.y, (Rr - Tr) >> 1, (Ri - Ti) >> 1, bX, bY);
  tw_mul_raw(w1.x, w1.y, (Sr + Ui) >> 1, (Si - Ur) >> 1, cX, cY);
  tw_mul_raw(w3.x, w3.y, (Sr - Ui) >> 1, (Si + Ur) >> 1, dX, dY);
}

// FIRST stage (arm_cfft_radix4_q15.c:192-330): inputs >>2, sums not halved except x[i0].
// The four operands are at most 8192 in magnitude after the shift, so no QADD16/QSUB16 of this
// stage can saturate; the plain adds below are exact.
ED_HD void bfly_first(int &ar, int &ai, int &br, int &bi, int &cr, int &ci, int &dr, int &di,
                      I2 w1, I2 w2, I2 w3) {
  ar >>= 2; ai >>= 2; br >>= 2; bi >>= 2; cr >>= 2; ci >>= 2; dr >>= 2; di >>= 2;
  int Rr = ar + cr, Ri = ai + ci, Sr = ar - cr, Si = ai - ci;
  int Tr = br + dr, Ti = bi + di, Ur = br - dr, Ui = bi - di;
  ar = (Rr + Tr) >> 1;
  ai = (Ri + Ti) >> 1;
  tw_mul(w2.x, w2.y, Rr - Tr, Ri - Ti, br, bi);
  tw_mul(w1.x, w1.y, Sr + Ui, Si - Ur, cr, ci);  // QSAX :287
  tw_mul(w3.x, w3.y, Sr - Ui, Si + Ur, dr, di);  // QASX :285
}

// LAST stage (arm_cfft_radix4_q15.c:492-557): no twiddles.
template <bool SAT>
ED_HD void bfly_last(int &ar, int &ai, int &br, int &bi, int &cr, int &ci, int &dr, int &di) {
  int Rr = qadd16<SAT>(ar, cr), Ri = qadd16<SAT>(ai, ci);
  int Sr = qsub16<SAT>(ar, cr), Si = qsub16<SAT>(ai, ci);
  int Tr = qadd16<SAT>(br, dr), Ti = qadd16<SAT>(bi, di);
  int Ur = qsub16<SAT>(br, dr), Ui = qsub16<SAT>(bi, di);
  ar = (Rr + Tr) >> 1; ai = (Ri + Ti) >> 1;
  br = (Rr - Tr) >> 1; bi = (Ri - Ti) >> 1;
  cr = (Sr + Ui) >> 1; ci = (Si - Ur) >> 1;  // SHSAX :540
  dr = (Sr - Ui) >> 1; di = (Si + Ur) >> 1;  // SHASX :544
}

// ------------------------------------------------------------------------------------------------
// arm_sqrt_q31 (lib/CMSIS/DSP/Source/FastMathFunctions/arm_sqrt_q31.c:50-135), exact.
// Three IEEE single-precision operations seed the Newton iteration: int->float (round to nearest),
// one multiply, one multiply, float->int (truncate).  None may be contracted into an FMA.
// ------------------------------------------------------------------------------------------------
#if defined(__CUDA_ARCH__)
ED_HD float ed_i2f(int v) { return __int2float_rn(v); }
ED_HD float ed_u2f(unsigned v) { return __uint2float_rn(v); }
ED_HD float ed_fmul(float a, float b) { return __fmul_rn(a, b); }
ED_HD int ed_f2i_rz(float v) { return __float2int_rz(v); }
ED_HD int ed_f_as_i(float v) { return __float_as_int(v); }
ED_HD float ed_i_as_f(int v) { return __int_as_float(v); }
ED_HD int ed_clz(int v) { return __clz(v); }
ED_HD float ed_sqrt_approx(float v) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }
#else
}  // namespace edison
#include <math.h>
#include <string.h>
namespace edison {
ED_HD float ed_i2f(int v) { volatile float f = (float)v; return f; }
ED_HD float ed_u2f(unsigned v) { volatile float f = (float)v; return f; }
ED_HD float ed_fmul(float a, float b) { volatile float f = a * b; return f; }
ED_HD int ed_f2i_rz(float v) { return (int)v; }
ED_HD int ed_f_as_i(float v) { int i; memcpy(&i, &v, 4); return i; }
ED_HD float ed_i_as_f(int v) { float f; memcpy(&f, &v, 4); return f; }
ED_HD int ed_clz(int v) { return v ? __builtin_clz((unsigned)v) : 32; }
ED_HD float ed_sqrt_approx(float v) { return sqrtf(v); }
#endif

#if defined(__CUDACC__)
#define ED_HD_NOINLINE static __host__ __device__ __noinline__
#else
#define ED_HD_NOINLINE inline
#endif

ED_HD int mul_q31(int a, int b) { return (int)(((long long)a * (long long)b) >> 31); }

// not inlined on the device: it is the rare path of mag_q15 and would bloat the hot loop
ED_HD_NOINLINE int sqrt_q31_newton(int in) {
  if (in <= 0) return 0;  // :127-133 (returns 0 with ARM_MATH_ARGUMENT_ERROR)
  int sign_bits = ed_clz(in) - 1;                      // :66
  int shift = sign_bits & ~1;                          // :69-76 (even shift)
  int number = (int)((unsigned)in << shift);           // :77
  int half = number >> 1;                              // :80
  float tf = ed_fmul(ed_i2f(number), 4.6566128731e-010f);  // :82
  int bits = 0x5f3759df - (ed_f_as_i(tf) >> 1);        // :86-87
  int var1 = ed_f2i_rz(ed_fmul(ed_i_as_f(bits), 1073741824.0f));  // :92
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int it = 0; it < 3; it++) {                     // :95-112
    int t = mul_q31(mul_q31(var1, var1), half);
    var1 = (int)((unsigned)mul_q31(var1, 0x30000000 - t) << 2);
  }
  var1 = (int)((unsigned)mul_q31(number, var1) << 1);  // :115
  return var1 >> (shift >> 1);                         // :118-125
}

// ------------------------------------------------------------------------------------------------
// cmpl_mag_q15 of one bin (firmware/src/audioprocessing.c:294-307):
//     mag = (q15)( arm_sqrt_q31(re*re + im*im) >> 16 )
// Exhaustive comparison over all 2^31 arguments (tools/sqrt_exceptions.md) shows that
//     arm_sqrt_q31(s) >> 16 == g := floor(sqrt(s / 2))
// except on 49,214 arguments, all of which sit next to a boundary of g:
//     g <  8192 : only s == 2 g^2 exactly, and then the result is g - 1 for a fixed set of g
//                 (a 8192-bit table, built at init from sqrt_q31_newton itself);
//     g >= 8192 : only when s - 2 g^2 <= 13 or 2 (g+1)^2 - s <= 11; those take the Newton path.
// So the common case costs one approximate square root and a one-sided integer correction, all on
// dlo = s - 2 g^2 (which the exception tests need anyway):  g -> g + 1  <=>  dlo >= 4 g + 2.
// ------------------------------------------------------------------------------------------------
constexpr int kSqrtExcWords = 8192 / 32;

// (1/sqrt 2)(1 - 2^-18) rounded to float.  y = sqrt.approx(float(s)) * kMagScale lies below sqrt(s/2) by a
// relative 2^-18 -- more than the three roundings and the approximation's error (< 2^-22 together) and
// less than 0.13 absolute (sqrt(s/2) <= 2^15) -- so trunc(y) is g or g - 1.
#define ED_MAG_SCALE 0.70710408687591552734f

// common case: g = floor(sqrt(s/2)), dlo = s - 2 g^2
ED_HD void mag_fast(unsigned s, int &g, int &dlo) {
  g = ed_f2i_rz(ed_fmul(ed_sqrt_approx(ed_u2f(s)), ED_MAG_SCALE));
  dlo = (int)(s - (unsigned)(2 * g) * (unsigned)g);
  const int t = 4 * g + 2;
  if (dlo >= t) { dlo -= t; g += 1; }
}
// the arguments on which arm_sqrt_q31 may differ from g (rare unless the spectrum is near empty)
ED_HD bool mag_is_rare(int g, int dlo) { return dlo == 0 || g >= 8192; }
ED_HD int mag_fix(unsigned s, int g, int dlo, const uint32_t *exc_bits) {
  if (g < 8192) {
    if (dlo == 0) g -= (int)((exc_bits[g >> 5] >> (g & 31)) & 1u);
  } else {
    const int dhi = 4 * g + 2 - dlo;  // 2 (g+1)^2 - s
    if (dlo <= 13 || dhi <= 11) g = (int)(int16_t)(sqrt_q31_newton((int)s) >> 16);
  }
  return g;
}

ED_HD int mag_from_sum(unsigned s, const uint32_t *exc_bits) {
  int g, dlo;
  mag_fast(s, g, dlo);
  if (mag_is_rare(g, dlo)) g = mag_fix(s, g, dlo, exc_bits);
  return g;
}

ED_HD int mag_q15(int re, int im, const uint32_t *exc_bits) {
  return mag_from_sum((unsigned)(re * re) + (unsigned)(im * im), exc_bits);
}
// two bins at once: both common paths run side by side, one branch covers the two rare tests
ED_HD void mag_q15_x2(int re0, int im0, int re1, int im1, const uint32_t *exc_bits, int &m0, int &m1) {
  const unsigned s0 = (unsigned)(re0 * re0) + (unsigned)(im0 * im0), s1 = (unsigned)(re1 * re1) + (unsigned)(im1 * im1);
  int d0, d1;
  mag_fast(s0, m0, d0);
  mag_fast(s1, m1, d1);
  if (mag_is_rare(m0, d0) | mag_is_rare(m1, d1)) {
    m0 = mag_fix(s0, m0, d0, exc_bits);
    m1 = mag_fix(s1, m1, d1, exc_bits);
  }
}

}  // namespace edison
