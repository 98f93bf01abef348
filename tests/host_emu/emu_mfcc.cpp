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

// mma.sync.m16n8k32.row.col.s32.u8.u8.s32 over the 32 lanes' fragments (PTX ISA "Matrix Fragments for mma.m16n8k32")
static void emu_mma_u8(int d[32][4], const I4 a[32], const uint32_t b0[32], const uint32_t b1[32]) {
  int A[16][32], B[32][8];
  for (int lane = 0; lane < 32; lane++) {
    const int g = lane >> 2, t = lane & 3;
    const uint32_t aw[4] = {(uint32_t)a[lane].x, (uint32_t)a[lane].y, (uint32_t)a[lane].z, (uint32_t)a[lane].w};
    for (int i = 0; i < 4; i++) {
      A[g][4 * t + i] = (aw[0] >> (8 * i)) & 255;
      A[g + 8][4 * t + i] = (aw[1] >> (8 * i)) & 255;
      A[g][16 + 4 * t + i] = (aw[2] >> (8 * i)) & 255;
      A[g + 8][16 + 4 * t + i] = (aw[3] >> (8 * i)) & 255;
      B[4 * t + i][g] = (b0[lane] >> (8 * i)) & 255;
      B[16 + 4 * t + i][g] = (b1[lane] >> (8 * i)) & 255;
    }
  }
  for (int lane = 0; lane < 32; lane++) {
    const int g = lane >> 2, t = lane & 3;
    for (int j = 0; j < 4; j++) {
      const int row = g + 8 * (j >> 1), col = 2 * t + (j & 1);
      int acc = 0;
      for (int k = 0; k < 32; k++) acc += A[row][k] * B[k][col];
      d[lane][j] += acc;
    }
  }
}

static void emu_mel_mma(int half, const uint32_t *spect, int *mel) {
  int p = 0;
  for (int mt = 0; mt < 2; mt++) {
    static int lo[32][4], hi[32][4];
    memset(lo, 0, sizeof lo); memset(hi, 0, sizeof hi);
    const int pend = p + (mt == 0 ? T.mel_n0 : T.mel_n1);
    for (; p < pend; p++) {
      const int ks = T.mel_ks[p];
      uint32_t b[4][32];
      for (int lane = 0; lane < 32; lane++)
        for (int k = 0; k < 4; k++) b[k][lane] = spect[mel_b_word(lane, ks, k >> 1, k & 1)];
      emu_mma_u8(lo, T.mel_a[p], b[0], b[1]);
      emu_mma_u8(hi, T.mel_a[p], b[2], b[3]);
    }
    for (int lane = 0; lane < 32; lane++)
      for (int j = 0; j < 4; j++) {
        const int fr = mel_c_frame(lane, j);
        if (fr < kSub) mel[(kSub * half + fr) * 32 + mel_c_filter(lane, mt, j)] = mel_finish(lo[lane][j], hi[lane][j]);
      }
  }
}

// force_sat: -1 = decide like the kernel, 0 = never, 1 = always (to cross-check the two paths)
extern "C" int emu_mfcc_batch(const int16_t *pcm, long n_frames, int16_t *mfcc, int8_t *net_in, int force_sat) {
  static uint32_t x[kXchWords], spect[kSpectWords];
  static int mel[kGroup * 32];
  const long n_groups = (n_frames + kGroup - 1) / kGroup;
  for (long grp = 0; grp < n_groups; grp++) {
    for (int g = 0; g < kGroup; g++) {
      const long f = grp * kGroup + g;
      if (f < n_frames) {
        const uint16_t *frame = (const uint16_t *)(pcm + f * kFrameLen);
        int peak = -1;
        static uint32_t w[32][16];
        for (int lane = 0; lane < 32; lane++) {
          for (int q = 0; q < 16; q++) w[lane][q] = frame[64 * q + 2 * lane] | ((uint32_t)frame[64 * q + 2 * lane + 1] << 16);
          int p = lane_peak(w[lane]);
          if (p > peak) peak = p;
        }
        for (int lane = 0; lane < 32; lane++)
          for (int e = 0; e < 2; e++) phase_a(lane, e, w[lane], T, x);
        bool sat = force_sat < 0 ? peak > kNoSatPeak : force_sat != 0;
        for (int lane = 0; lane < 32; lane++)
          for (int e = 0; e < 2; e++) { if (sat) phase_b<true>(lane, e, T, x); else phase_b<false>(lane, e, T, x); }
        uint32_t *row = spect + (g & (kSub - 1)) * kSpectRowWords;
        for (int lane = 0; lane < 32; lane++)
          for (int s = 0; s < 2; s++) { if (sat) phase_c<true>(lane, s, T, x, row); else phase_c<false>(lane, s, T, x, row); }
      }
      if ((g & (kSub - 1)) == kSub - 1) emu_mel_mma(g / kSub, spect, mel);
    }
    I2 *z = reinterpret_cast<I2 *>(x);
    int *out = reinterpret_cast<int *>(x) + 2 * kGroup * 16;
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
