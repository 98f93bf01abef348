// mfcc_q15_kernel.cu -- K1: batched q15 MFCC front-end for sm_100a.
//
// Replaces audioCalcMFCCs (firmware/src/audioprocessing.c:116-215) + mfccToNetInput
// (firmware/src/app.c:675-695) over a batch of frames.  One warp owns one frame at a time and
// walks groups of 8 consecutive frames; the arithmetic lives in mfcc_q15_core.cuh.
//
// HBM traffic per frame: 2,048 B of PCM read once (coalesced 128 B per warp-load), 64 B of q15
// coefficients and/or 13 B of int8 net input written.  Everything else stays in registers and
// shared memory.  The kernel is integer-issue bound, so the grid is persistent:
// one 20-warp CTA per SM.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "edison_internal.h"
#include "../../include/edison_b200.h"

namespace edison {

constexpr int kWarpsPerCta = 20;
constexpr int kWindowStride = EDISON_NET_IN;   // STRIDED instantiations: one 13-byte row into each stream's 403-byte window (stream.cu)
constexpr int kThreads = kWarpsPerCta * 32;

struct WarpSmem {
  uint32_t x[kXchWords];        // exchange buffer (packed complex); doubles as DCT scratch at group end
  uint32_t spect[kSpectWords];  // byte planes of the magnitude spectrum, 4 frames
  int mel[kGroup * 32];
};
// the DCT scratch (z: 8*16 I2 = 1 KiB, out: 8*32 int = 1 KiB) aliases x, which is idle by then
static_assert(sizeof(I2) * kGroup * 16 + sizeof(int) * kGroup * 32 <= sizeof(uint32_t) * kXchWords, "dct scratch");

struct CtaSmem {
  MfccTables T;
  WarpSmem w[kWarpsPerCta];
};

__device__ __forceinline__ uint32_t ld_nc_u32(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.b32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t ld_nc_u16(const uint16_t *p) {
  uint16_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(v) : "l"(p));
  return v;
}

// the 16 words of frame f this lane owns: w[q] = samples 64q + 2 lane, 64q + 2 lane + 1
// ALIGNED: every frame starts on a 4-byte boundary (pcm 4-byte aligned, even clip stride); decided at launch
template <bool ALIGNED>
__device__ __forceinline__ void load_frame(const MfccLaunch &P, long long f, int lane, uint32_t (&w)[16]) {
  const long long clip = f / P.frames_per_clip;
  const int fic = (int)(f - clip * P.frames_per_clip);
  const int16_t *frame = P.pcm + clip * P.clip_stride + (long long)fic * kFrameLen;
  if (ALIGNED) {
    const uint32_t *fw = reinterpret_cast<const uint32_t *>(frame);
#pragma unroll
    for (int q = 0; q < 16; q++) w[q] = ld_nc_u32(fw + 32 * q + lane);
  } else {  // odd sample offset (clip stride not a multiple of 2 samples)
    const uint16_t *fh = reinterpret_cast<const uint16_t *>(frame);
#pragma unroll
    for (int q = 0; q < 16; q++)
      w[q] = ld_nc_u16(fh + 64 * q + 2 * lane) | (ld_nc_u16(fh + 64 * q + 2 * lane + 1) << 16);
  }
}

__device__ __forceinline__ void mma_u8(int (&d)[4], const I4 &a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
               : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b0), "r"(b1));
}

// mel filterbank of the 4 frames whose magnitude planes sit in `spect` -> mel[(4 half + frame) * 32 + filter]
__device__ __forceinline__ void mel_mma(int lane, int half, const MfccTables &T, const uint32_t *spect, int *mel) {
  int p = 0;
#pragma unroll
  for (int mt = 0; mt < 2; mt++) {
    int lo[4] = {0, 0, 0, 0}, hi[4] = {0, 0, 0, 0};
    const int pend = p + (mt == 0 ? T.mel_n0 : T.mel_n1);
#pragma unroll 2
    for (; p < pend; p++) {
      const int ks = T.mel_ks[p];
      const I4 a = T.mel_a[p][lane];
      mma_u8(lo, a, spect[mel_b_word(lane, ks, 0, 0)], spect[mel_b_word(lane, ks, 0, 1)]);
      mma_u8(hi, a, spect[mel_b_word(lane, ks, 1, 0)], spect[mel_b_word(lane, ks, 1, 1)]);
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int fr = mel_c_frame(lane, j);
      if (fr < kSub) mel[(kSub * half + fr) * 32 + mel_c_filter(lane, mt, j)] = mel_finish(lo[j], hi[j]);
    }
  }
}

// dct2_q15 + outputs for the 8 frames of group grp, whose mel spectra sit in W.mel
// SMALL: a group holds only `gf` (1, 2 or 4) frames -- f0 = grp * gf, slots gf..7 carry stale data and are not stored
template <bool STRIDED, bool SMALL>
__device__ __forceinline__ void group_tail(const MfccLaunch &P, const MfccTables &T, WarpSmem &W, int lane, long long grp,
                                           long long n_frames, int gf) {
  // dct2_q15 for the 8 frames of the group, 4 lanes per frame
  I2 *z = reinterpret_cast<I2 *>(W.x);
  int *out = reinterpret_cast<int *>(W.x) + 2 * kGroup * 16;
  dct_stage1(lane, T, W.mel, z);
  __syncwarp();
  dct_stage2(lane, z);
  __syncwarp();
#pragma unroll
  for (int it = 0; it < 4; it++) dct_split(32 * it + lane, T, z, out);
  __syncwarp();
  const long long f0 = grp * (SMALL ? gf : kGroup);
  if (P.mfcc) {  // 8 frames x 64 B, one 16-byte store per lane
    const int fr = lane >> 2, part = lane & 3;
    if (f0 + fr < n_frames && (!SMALL || fr < gf)) {
      const int *o = out + 32 * fr + 8 * part;
      uint4 v;
      v.x = (uint32_t)(o[0] & 0xffff) | ((uint32_t)o[1] << 16);
      v.y = (uint32_t)(o[2] & 0xffff) | ((uint32_t)o[3] << 16);
      v.z = (uint32_t)(o[4] & 0xffff) | ((uint32_t)o[5] << 16);
      v.w = (uint32_t)(o[6] & 0xffff) | ((uint32_t)o[7] << 16);
      *reinterpret_cast<uint4 *>(P.mfcc + (f0 + fr) * 32 + 8 * part) = v;
    }
  }
  if (P.net_in) {  // 8 frames x 13 int8 (mfccToNetInput)
    for (int idx = lane; idx < kGroup * kNCoeff; idx += 32) {
      const int fr = idx / kNCoeff, c = idx - fr * kNCoeff;
            // STRIDED (the streaming path: one row of each stream's window) is a separate instantiation: with the stride as a run-time
      // value in the batch kernel, bench.py's launch went from 10.87 to 11.29 ms (register allocation of the hot loop shifts)
      if (f0 + fr < n_frames && (!SMALL || fr < gf))
        P.net_in[(f0 + fr) * (long long)(STRIDED ? kWindowStride : kNCoeff) + c] = (int8_t)net_input_q7(out[32 * fr + c], T);
    }
  }
  __syncwarp();
}

// SMALL (latency path, launches of < 11,841 frames): groups of P.group_frames = 1, 2 or 4 frames instead of 8, so that a hop of a
// thousand streams is one frame on each of a thousand warps spread over all SMs instead of eight frames in a row on 128 warps.
// The batch instantiations do not read P.group_frames: their hot loop is unchanged.
template <bool ALIGNED, bool STRIDED, bool SMALL>
__global__ void __launch_bounds__(kThreads, 1) mfcc_q15_kernel(MfccLaunch P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  CtaSmem &S = *reinterpret_cast<CtaSmem *>(smem_raw);
  {  // tables: global -> shared, once per CTA
    const int4 *src = reinterpret_cast<const int4 *>(P.tables);
    int4 *dst = reinterpret_cast<int4 *>(&S.T);
    for (int i = threadIdx.x; i < (int)(sizeof(MfccTables) / 16); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const MfccTables &T = S.T;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  WarpSmem &W = S.w[warp];
  const long long n_frames = P.n_clips * P.frames_per_clip;
  const int gf = SMALL ? P.group_frames : kGroup;   // frames per group
  const long long n_groups = (n_frames + gf - 1) / gf;
  const int warps_per_cta = blockDim.x >> 5;   // a multiple of 4, at most kWarpsPerCta (small launches run fewer warps)
  // work is dealt to barrier groups (4 warps) CTA-minor: the last, partial trip then leaves a group or two busy on EVERY SM
  // instead of filling the first few SMs completely (a 100,000-frame hop of the streaming path is 4.2 trips)
  const long long warp_global = ((long long)(warp >> 2) * gridDim.x + blockIdx.x) * 4 + (warp & 3);
  const long long warp_stride = (long long)gridDim.x * warps_per_cta;

  uint32_t w[16];
#pragma unroll
  for (int q = 0; q < 16; q++) w[q] = 0;
  if (warp_global < n_groups) load_frame<ALIGNED>(P, warp_global * gf, lane, w);
  // Lockstep in groups of four warps, one warp per scheduler (named barrier 1 + warp/4, once per frame).  Warps that run
  // the same code at the same time share their instruction fetches (free-running warps: 12.2 ms for bench.py's launch), but
  // warps that share a SCHEDULER then also want the same pipe at the same time -- the FFT stages are ALU/IMAD throughput
  // bound, the magnitude code latency bound -- and a CTA-wide barrier (11.9 ms) gives up that mixing; groups that drift
  // against each other keep both (10.9 ms).  profiles/r01k.
  // The four warps of a group run the same trip count and compute every frame of their trips, frames past the end from
  // stale registers (no address depends on data; nothing of theirs is stored): the hot loop has no "valid" branches.
  const int bar_id = 1 + (warp >> 2);   // groups of 8, 10, 12 warps measured 11.0, 11.3, 11.35 ms
  const long long n_iter = (n_groups + warp_stride - 1) / warp_stride;
  for (long long it = 0, grp = warp_global; it < n_iter; it++, grp += warp_stride) {
    if (grp - (warp & 3) >= n_groups) break;   // none of the four warps of this barrier group has work left: all four leave
#pragma unroll 1
    for (int g = 0; g < gf; g++) {
      const long long f = grp * gf + g;
      uint32_t *row = W.spect + (g & (kSub - 1)) * kSpectRowWords;
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id));
      const bool sat = __reduce_max_sync(0xffffffffu, lane_peak(w)) > kNoSatPeak;
#pragma unroll 1
      for (int e = 0; e < 2; e++) phase_a(lane, e, w, T, W.x);
      // prefetch the next frame this warp will process; it lands while phases B, C and mel run
      const long long fn = g + 1 < gf ? f + 1 : (grp + warp_stride) * gf;
      if (fn < n_frames) load_frame<ALIGNED>(P, fn, lane, w);
      __syncwarp();
      if (sat) {
#pragma unroll 1
        for (int e = 0; e < 2; e++) phase_b<true>(lane, e, T, W.x);
      } else {
#pragma unroll 1
        for (int e = 0; e < 2; e++) phase_b<false>(lane, e, T, W.x);
      }
      __syncwarp();
      if (sat) {
#pragma unroll 1
        for (int s = 0; s < 2; s++) phase_c<true>(lane, s, T, W.x, row);
      } else {
#pragma unroll 1
        for (int s = 0; s < 2; s++) phase_c<false>(lane, s, T, W.x, row);
      }
      __syncwarp();
      if ((g & (kSub - 1)) == kSub - 1 || (SMALL && g == gf - 1)) {
        mel_mma(lane, g / kSub, T, W.spect, W.mel);
        __syncwarp();
      }
    }
    group_tail<STRIDED, SMALL>(P, T, W, lane, grp, n_frames, gf);
  }
}

template <bool A, bool S, bool M> static cudaError_t configure_one(void) {
  return cudaFuncSetAttribute(mfcc_q15_kernel<A, S, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(CtaSmem));
}
cudaError_t mfcc_q15_configure(void) {
  cudaError_t e = configure_one<true, false, false>();
  if (e == cudaSuccess) e = configure_one<false, false, false>();
  if (e == cudaSuccess) e = configure_one<true, true, false>();
  if (e == cudaSuccess) e = configure_one<false, true, false>();
  if (e == cudaSuccess) e = configure_one<true, false, true>();
  if (e == cudaSuccess) e = configure_one<false, false, true>();
  if (e == cudaSuccess) e = configure_one<true, true, true>();
  if (e == cudaSuccess) e = configure_one<false, true, true>();
  return e;
}

template <bool A, bool S> static void launch_one(const MfccLaunch &p, bool small, unsigned ctas, unsigned threads, cudaStream_t stream) {
  if (small) mfcc_q15_kernel<A, S, true><<<ctas, threads, sizeof(CtaSmem), stream>>>(p);
  else mfcc_q15_kernel<A, S, false><<<ctas, threads, sizeof(CtaSmem), stream>>>(p);
}

cudaError_t launch_mfcc_q15(const MfccLaunch &p_in, int sm_count, cudaStream_t stream) {
  MfccLaunch p = p_in;
  const long long n_frames = p.n_clips * p.frames_per_clip;
  if (n_frames <= 0) return cudaSuccess;
  // frames per group: 8, or -- when that would leave warp slots of the machine idle -- the smallest of 1, 2, 4 whose groups still
  // fit the machine's warps in one trip (a small launch is a latency problem: frames in parallel on many warps, not in a row on few)
  const long long slots = (long long)sm_count * kWarpsPerCta;
  int gf = kGroup;
  for (int c = 1; c < kGroup; c *= 2)
    if ((n_frames + c - 1) / c <= slots) { gf = c; break; }
  p.group_frames = gf;
  const long long n_groups = (n_frames + gf - 1) / gf;
  // barrier groups of four warps are dealt across the SMs first: CTAs = min(SMs, groups of four), then as many warps per CTA as needed
  long long ctas = (n_groups + 3) / 4;
  if (ctas > sm_count) ctas = sm_count;
  long long warps = (n_groups + ctas - 1) / ctas;
  warps = warps >= kWarpsPerCta ? kWarpsPerCta : (warps + 3) / 4 * 4;
  const unsigned threads = (unsigned)warps * 32;
  const bool aligned = (reinterpret_cast<uintptr_t>(p.pcm) & 3u) == 0 && (p.clip_stride & 1) == 0;
  if (p.net_in && p.net_in_stride != kNCoeff && p.net_in_stride != kWindowStride) return cudaErrorInvalidValue;   // packed rows or window rows only
  const bool strided = p.net_in && p.net_in_stride == kWindowStride;
  const bool small = gf != kGroup;
  if (aligned && !strided) launch_one<true, false>(p, small, (unsigned)ctas, threads, stream);
  else if (!strided) launch_one<false, false>(p, small, (unsigned)ctas, threads, stream);
  else if (aligned) launch_one<true, true>(p, small, (unsigned)ctas, threads, stream);
  else launch_one<false, true>(p, small, (unsigned)ctas, threads, stream);
  count_launch();
  return cudaGetLastError();
}

// ---- debug kernels (device arithmetic against itself; include/edison_b200.h section 4) ---------
__global__ void debug_sqrt_kernel(const int32_t *in, size_t n, int32_t *out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = sqrt_q31_newton(in[i]);
}
cudaError_t launch_debug_sqrt(const int32_t *d_in, size_t n, int32_t *d_out, cudaStream_t stream) {
  if (!n) return cudaSuccess;
  debug_sqrt_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(d_in, n, d_out);
  count_launch();
  return cudaGetLastError();
}

__global__ void debug_mag_sweep_kernel(uint32_t lo, uint32_t hi, const uint32_t *exc, unsigned long long *out2) {
  unsigned long long mism = 0, sum = 0;
  for (unsigned long long s = (unsigned long long)lo + (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; s < hi;
       s += (unsigned long long)gridDim.x * blockDim.x) {
    const int exact = (int)(int16_t)(sqrt_q31_newton((int)(unsigned)s) >> 16);
    const int fast = mag_from_sum((unsigned)s, exc);
    mism += (exact != fast);
    sum += (unsigned long long)(unsigned)exact * ((s & 1023ull) + 1ull);
  }
  for (int o = 16; o; o >>= 1) {
    mism += __shfl_xor_sync(0xffffffffu, mism, o);
    sum += __shfl_xor_sync(0xffffffffu, sum, o);
  }
  if ((threadIdx.x & 31) == 0) { atomicAdd(out2, mism); atomicAdd(out2 + 1, sum); }
}
cudaError_t launch_debug_mag_sweep(uint32_t lo, uint32_t hi, const uint32_t *d_exc, unsigned long long *d_out2,
                                   cudaStream_t stream) {
  debug_mag_sweep_kernel<<<148 * 8, 256, 0, stream>>>(lo, hi, d_exc, d_out2);
  count_launch();
  return cudaGetLastError();
}

}  // namespace edison
