// emu_mfcc.cpp -- host emulation of the K1 warp program (TEST INFRASTRUCTURE).
//
// Compiles edison_b200/csrc/mfcc_q15_core.cuh for the host and runs the kernel's phase sequence
// (edison_b200/csrc/mfcc_q15_kernel.cu) with the 32 lanes executed one after another between the
// points where the kernel has a __syncwarp().  It exists so that the index arithmetic of the
// CUDA source (exchange layouts, twiddle tables, bit reversal, DCT batching) can be checked
// against the oracle in a container without a GPU.  It is not part of libedison_b200.so and is
// never used to produce results: the -m gpu parity tests run the real kernel.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "edison_internal.h"

using namespace edison;

static MfccTables T;

extern "C" int emu_init(const char *mel_blob) {
  char err[256] = "";
  int r = build_mfcc_tables(&T, mel_blob, err, sizeof err);
  if (r) fprintf(stderr, "emu_init: %s\n", err);
  return r;
}

// force_sat: -1 = decide like the kernel, 0 = never, 1 = always (to cross-check the two paths)
extern "C" int emu_mfcc_batch(const int16_t *pcm, long n_frames, int16_t *mfcc, int8_t *net_in, int force_sat) {
  static int xr[kXchPlane], xi[kXchPlane], spect[512], mel[kGroup * 32];
  const long n_groups = (n_frames + kGroup - 1) / kGroup;
  for (long grp = 0; grp < n_groups; grp++) {
    for (int g = 0; g < kGroup; g++) {
      const long f = grp * kGroup + g;
      if (f >= n_frames) { for (int l = 0; l < 32; l++) mel[32 * g + l] = 0; continue; }
      const uint16_t *frame = (const uint16_t *)(pcm + f * kFrameLen);
      int peak = -1;
      uint32_t w[32][16];
      for (int lane = 0; lane < 32; lane++) {
        for (int q = 0; q < 16; q++) w[lane][q] = frame[64 * q + 2 * lane] | ((uint32_t)frame[64 * q + 2 * lane + 1] << 16);
        int p = lane_peak(w[lane]);
        if (p > peak) peak = p;
      }
      for (int lane = 0; lane < 32; lane++)
        for (int e = 0; e < 2; e++) phase_a(lane, e, w[lane], T, xr, xi);
      bool sat = force_sat < 0 ? peak > kNoSatPeak : force_sat != 0;
      for (int lane = 0; lane < 32; lane++)
        for (int e = 0; e < 2; e++) { if (sat) phase_b<true>(lane, e, T, xr, xi); else phase_b<false>(lane, e, T, xr, xi); }
      for (int lane = 0; lane < 32; lane++)
        for (int i = 0; i < 8; i++) { if (sat) phase_c<true>(lane, i, T, xr, xi, spect); else phase_c<false>(lane, i, T, xr, xi, spect); }
      for (int lane = 0; lane < 32; lane++) mel[32 * g + lane] = phase_mel(lane, T, spect);
    }
    I2 *z = reinterpret_cast<I2 *>(xr);
    int *out = xr + 2 * kGroup * 16;
    for (int lane = 0; lane < 32; lane++) dct_stage1(lane, T, mel, z);
    for (int lane = 0; lane < 32; lane++) dct_stage2(lane, z);
    for (int it = 0; it < 4; it++)
      for (int lane = 0; lane < 32; lane++) dct_split(32 * it + lane, T, z, out);
    for (int fr = 0; fr < kGroup; fr++) {
      const long f = grp * kGroup + fr;
      if (f >= n_frames) break;
      if (mfcc) for (int k = 0; k < 32; k++) mfcc[f * 32 + k] = (int16_t)out[32 * fr + k];
      if (net_in) for (int c = 0; c < kNCoeff; c++) net_in[f * kNCoeff + c] = (int8_t)net_input_q7(out[32 * fr + c], T);
    }
  }
  return 0;
}

extern "C" int emu_mag_q15(int re, int im) { return mag_q15(re, im, T.sqrt_exc); }
extern "C" int emu_sqrt_q31(int in) { return sqrt_q31_newton(in); }
