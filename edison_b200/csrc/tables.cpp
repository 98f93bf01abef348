// tables.cpp -- constant tables of the q15 front-end, generated on the host at edison_init().
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "edison_internal.h"
#include "../../include/edison_b200.h"

namespace edison {

// ------------------------------------------------------------------------------------------------
// Tables.  arm_common_tables.c is missing from the reference tree (.MISSING_LARGE_BLOBS:4); the
// formulas below reproduce the data inside firmware/src/lib/CMSIS/Lib/GCC/libarm_cortexM4lf_math.a
// bit for bit (twiddleCoef_N_q15 = sat16(floor(cos|sin * 32768)); realCoefA/BQ15 =
// floor(x * 32768 + 0.5)).  FNV-1a checksums of the generated tables are compared with the
// checksums of the reference data so that a libm with different rounding cannot go unnoticed.
// ------------------------------------------------------------------------------------------------
static int16_t q15_floor_sat(double v) {
  double f = floor(v);
  if (f > 32767.0) f = 32767.0;
  if (f < -32768.0) f = -32768.0;
  return (int16_t)f;
}
static uint64_t fnv1a(const void *p, size_t n, uint64_t h = 1469598103934665603ull) {
  const unsigned char *b = (const unsigned char *)p;
  for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
  return h;
}

int build_mfcc_tables(MfccTables *Tp, const char *mel_blob_path, char *err, size_t errlen) {
  MfccTables &T = *Tp;
  memset(&T, 0, sizeof T);
  const double two_pi = 6.283185307179586476925286766559;
  static I2 tw1024[768];
  int16_t flat[1536];
  for (int i = 0; i < 768; i++) {
    flat[2 * i] = q15_floor_sat(cos(two_pi * i / 1024.0) * 32768.0);
    flat[2 * i + 1] = q15_floor_sat(sin(two_pi * i / 1024.0) * 32768.0);
    tw1024[i] = I2{flat[2 * i], flat[2 * i + 1]};
  }
  const uint64_t h1024 = fnv1a(flat, sizeof flat);
  int16_t flat16[24];
  for (int i = 0; i < 12; i++) {
    flat16[2 * i] = q15_floor_sat(cos(two_pi * i / 16.0) * 32768.0);
    flat16[2 * i + 1] = q15_floor_sat(sin(two_pi * i / 16.0) * 32768.0);
    T.tw16[i] = I2{flat16[2 * i], flat16[2 * i + 1]};
  }
  int16_t flatAB[64];
  for (int n = 0; n < 16; n++) {
    const double th = two_pi * (256.0 * n) / 8192.0;  // stride 2*256 q15 in an 8192-entry table
    flatAB[2 * n] = q15_floor_sat(32768.0 * 0.5 * (1.0 - sin(th)) + 0.5);
    flatAB[2 * n + 1] = q15_floor_sat(32768.0 * 0.5 * (-cos(th)) + 0.5);
    flatAB[32 + 2 * n] = q15_floor_sat(32768.0 * 0.5 * (1.0 + sin(th)) + 0.5);
    flatAB[32 + 2 * n + 1] = q15_floor_sat(32768.0 * 0.5 * (cos(th)) + 0.5);
    T.realA[n] = I2{flatAB[2 * n], flatAB[2 * n + 1]};
    T.realB[n] = I2{flatAB[32 + 2 * n], flatAB[32 + 2 * n + 1]};
  }
  const uint64_t h16 = fnv1a(flat16, sizeof flat16), hAB = fnv1a(flatAB, sizeof flatAB);
  // checksums of the same tables extracted from the reference's prebuilt CMSIS library
  // (tools/table_checksums.py prints them)
  const uint64_t kH1024 = 0xe5a1b5649a0f9559ull, kH16 = 0x10feb6fa3ba94019ull, kHAB = 0xf44c1337e7aa76fdull;
  if (h1024 != kH1024 || h16 != kH16 || hAB != kHAB) {
    snprintf(err, errlen, "generated q15 tables differ from the CMSIS reference data (libm rounding?)");
    return EDISON_E_DATA;
  }
  for (int k = 0; k < 3; k++)
    for (int e = 0; e < 2; e++)
      for (int lane = 0; lane < 32; lane++) {
        for (int qp = 0; qp < 4; qp++) T.tw1[k][e][qp][lane] = tw1024[(k + 1) * (64 * qp + 2 * lane + e)];
        T.tw2[k][e][lane] = tw1024[(k + 1) * 4 * (2 * lane + e)];
      }
  for (int j = 0; j < 16; j++)
    for (int k = 0; k < 3; k++) T.tw3[k][j] = tw1024[(k + 1) * 16 * j];
  for (int j = 0; j < 4; j++)
    for (int k = 0; k < 3; k++) T.tw4[k][j] = tw1024[(k + 1) * 64 * j];
  // sqrt exception bits (q15_arith.cuh: mag_q15): bit g set when arm_sqrt_q31(2 g^2) >> 16 == g - 1
  for (int g = 0; g < 8192; g++) {
    const int f = sqrt_q31_newton(2 * g * g) >> 16;
    if (f != g && f != g - 1) { snprintf(err, errlen, "sqrt table: unexpected value at g=%d", g); return EDISON_E_DATA; }
    if (f == g - 1) T.sqrt_exc[g >> 5] |= 1u << (g & 31);
  }
  // mel blob (tools/convert_mel_constants.py)
  FILE *fh = fopen(mel_blob_path, "rb");
  if (!fh) { snprintf(err, errlen, "cannot open %s", mel_blob_path); return EDISON_E_DATA; }
  unsigned char hdr[24];
  uint32_t v[4];
  if (fread(hdr, 1, 24, fh) != 24 || memcmp(hdr, "EDMELQ15", 8)) { fclose(fh); snprintf(err, errlen, "%s: bad header", mel_blob_path); return EDISON_E_DATA; }
  memcpy(v, hdr + 8, 16);
  if (v[0] != 1 || v[1] != (uint32_t)kNMel || v[2] > (uint32_t)kMelMaxW || v[3] != 128) { fclose(fh); snprintf(err, errlen, "%s: unsupported mel table", mel_blob_path); return EDISON_E_DATA; }
  static int16_t mel_start[kNMel], mel_cnt[kNMel], mel_w[kMelMaxW];
  bool ok = fread(mel_start, 2, kNMel, fh) == (size_t)kNMel && fread(mel_cnt, 2, kNMel, fh) == (size_t)kNMel &&
            fread(mel_w, 2, v[2], fh) == v[2];
  fclose(fh);
  if (!ok) { snprintf(err, errlen, "%s: truncated", mel_blob_path); return EDISON_E_DATA; }
  // dense [filter][bin] u8 matrix from the compact form, then the non-zero (16 filters x 32 bins) tiles as
  // mma.m16n8k32 A fragments: lane (g, t) holds a0 = W[16mt+g][32ks+4t..+3], a1 = W[16mt+g+8][same],
  // a2 = W[16mt+g][32ks+16+4t..+3], a3 = W[16mt+g+8][same]
  static uint8_t dense[kNMel][512];
  memset(dense, 0, sizeof dense);
  int off = 0;
  for (int m = 0; m < kNMel; m++) {
    if (mel_start[m] < 0 || mel_cnt[m] < 0 || mel_start[m] + mel_cnt[m] > 512) { snprintf(err, errlen, "mel filter %d reaches past bin 511", m); return EDISON_E_DATA; }
    for (int i = 0; i < mel_cnt[m]; i++) {
      const int wv = mel_w[off + i];
      if (wv < 0 || wv > 255) { snprintf(err, errlen, "mel weight %d of filter %d does not fit u8", wv, m); return EDISON_E_DATA; }
      dense[m][mel_start[m] + i] = (uint8_t)wv;
    }
    off += mel_cnt[m];
  }
  if (off != (int)v[2]) { snprintf(err, errlen, "mel counts do not sum to the weight count"); return EDISON_E_DATA; }
  int np = 0;
  T.mel_n0 = T.mel_n1 = 0;
  for (int mt = 0; mt < 2; mt++)
    for (int ks = 0; ks < 16; ks++) {
      bool any = false;
      for (int m = 0; m < 16 && !any; m++)
        for (int k = 0; k < 32; k++) any |= dense[16 * mt + m][32 * ks + k] != 0;
      if (!any) continue;
      for (int lane = 0; lane < 32; lane++) {
        const int g = lane >> 2, t = lane & 3;
        auto word = [&](int row, int k0) {
          uint32_t w = 0;
          for (int i = 0; i < 4; i++) w |= (uint32_t)dense[16 * mt + row][32 * ks + k0 + i] << (8 * i);
          return (int)w;
        };
        T.mel_a[np][lane] = I4{word(g, 4 * t), word(g + 8, 4 * t), word(g, 16 + 4 * t), word(g + 8, 16 + 4 * t)};
      }
      T.mel_ks[np] = (uint8_t)ks;
      (mt == 0 ? T.mel_n0 : T.mel_n1)++;
      np++;
    }
  T.in_scale = 1; T.in_min = -128; T.in_max = 127;  // overwritten from the model blob's trailer
  return EDISON_OK;
}

}  // namespace edison
