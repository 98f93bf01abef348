// nnom_q7_umma.cu -- K2 on the Blackwell tensor path: the int8 NNoM keyword-spotting net
// (firmware/src/ai/nnom/kws_nnom/weights.h:138-161; model_run, firmware/src/ai/nnom/src/core/nnom.c:996-1041) as
// tcgen05.mma.kind::i8 implicit GEMMs with accumulators in TMEM.  The plan -- shared-memory layout, MMA list, unit order,
// epilogues -- is k2_umma_plan.h (also run on the host by tests/host_emu/emu_k2.cpp); this file is its executor:
//
//   * persistent grid, one CTA per SM, 32 clips per batch, activations of a batch never leave shared memory;
//   * warp 16 (converged, one lane issuing): waits for the TMEM slot of the next unit (mbarrier empty[slot]), issues the unit's MMAs from
//     shared-memory descriptors (operands are never copied: a filter tap is a shifted start address), commits them to
//     mbarrier full[slot]; it also starts the bulk copies (cp.async.bulk -> SASS UBLKCP) of the weight image and of each
//     batch's 32 x 403 input bytes;
//   * warps 0..15, four teams of four (team t drains slot t): wait on full[slot], read their lane quarter of the accumulators with tcgen05.ld (lane = clip), run the
//     integer epilogue of arm_convolve_HWC_q7_fast_nonsquare.c / nnom_local.c (bias seed, >> shift, SSAT8, ReLU, 2x1 max
//     pool), store 16 channels as one 16-byte word in the next layer's operand layout, make the stores visible to the
//     tensor core's asynchronous proxy, and release the slot (mbarrier empty[slot]);
//   * the unit order interleaves the head of batch b with the tail of batch b - 1 so that every unit's producers are
//     at least four units older: the only waits are those of the four-slot TMEM ring (one team of four epilogue warps per slot).
// Two kernels share this machinery: nnom_q7_umma_kernel (the whole window, 21 units per iteration) and nnom_q7_umma_incr_kernel (the streaming
// path's incremental mode: the bottom row of every layer per hop on an eight-unit schedule, older rows from a row cache in HBM; k2_umma_plan.h).
// There is no CPU path and no legacy mma.sync path behind these entry points.
#include <stdio.h>
#include <stdlib.h>

#include "edison_internal.h"
#include "../../include/edison_b200.h"
#include "k2_umma_plan.h"

namespace edison {
using namespace k2;

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// The issuing warp runs converged; every instruction that one thread issues on behalf of the CTA is predicated on elect.sync
// inside the same asm block (the form ptxas recognises as single-thread issue: ELECT + predicated UTCIMMA, no divergence loop).
// elect.sync over the full mask names the same lane every time, which tcgen05.commit needs (it tracks the MMAs of ITS thread).
__device__ __forceinline__ void mma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p, e;\n\t.reg .b32 r;\n\telect.sync r|e, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "@e tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n"
               :: "r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("{\n\t.reg .pred e;\n\t.reg .b32 r;\n\telect.sync r|e, 0xffffffff;\n\t"
               "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}\n" :: "r"(bar) : "memory");
}
__device__ __forceinline__ bool elect_one() {
  uint32_t e;
  asm volatile("{\n\t.reg .pred e;\n\t.reg .b32 r;\n\telect.sync r|e, 0xffffffff;\n\tselp.u32 %0, 1, 0, e;\n\t}\n" : "=r"(e));
  return e != 0;
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
               "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
               "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}\n"
               :: "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}\n" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}\n" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

struct DevCtx {
  uint8_t *smem;
  uint32_t taddr;   // TMEM address of this thread's lane quarter at the unit's slot
  int lane, q, g;
  long long clip0;
  int nclips;
  int8_t *logits, *softmax;
  int softmax_mode;
  int *hist;        // shared-memory class histogram
  bool counting;    // the caller asked for the class histogram
  float *filt;      // streaming post-processing (null: none)
  int32_t *detect;
  uint8_t *cache_batch;               // incremental mode: this unit's batch in the row cache
  uint32_t cache_w_c2, cache_w_c3;    //   offsets of the slots this hop writes (kNoCache: no caching)
  long long t_ld;   // TRACE builds: clock64() when the last tcgen05.ld had landed
  bool tracing;
  __device__ __forceinline__ void ld16(uint32_t col, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr + col));
  }
  __device__ __forceinline__ void ld_wait() {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    if (tracing) t_ld = clock64();
  }
  __device__ __forceinline__ void st16(uint32_t off, Vec16 v) { *reinterpret_cast<uint4 *>(smem + off) = make_uint4(v.x, v.y, v.z, v.w); }
  __device__ __forceinline__ void st16_cache(uint32_t off, Vec16 v) { *reinterpret_cast<uint4 *>(cache_batch + off) = make_uint4(v.x, v.y, v.z, v.w); }
  __device__ __forceinline__ int par(int i) const { return reinterpret_cast<const int *>(smem + OFF_PAR)[i]; }
  __device__ __forceinline__ Int4 par4(int i) const {
    const int4 v = *reinterpret_cast<const int4 *>(smem + OFF_PAR + 4 * i);
    return Int4{v.x, v.y, v.z, v.w};
  }
  __device__ __forceinline__ void count(int cls) { atomicAdd(&hist[cls], 1); }
  // app.c:336-350: netOutFilt = 0.9 netOutFilt + (1.0 - 0.9) netOutFloat, evaluated in double (the literals are doubles, the
  // firmware has no -fsingle-precision-constant) and stored as float; arm_max_f32 (first maximum); spotted when > 0.5.
  // The ten state loads are issued by prefetch() before the dense unit's accumulators are read: inside the loop each sat behind the
  // previous value's store, ten global-memory latencies in a row, and in the incremental mode the dense unit's team is the critical path.
  float pf[10];
  __device__ __forceinline__ void prefetch(long long clip) {
#pragma unroll
    for (int i = 0; i < 10; i++) pf[i] = filt[clip * 10 + i];
  }
  __device__ __forceinline__ void post(long long clip, const int8_t *sm) {
    const double alpha = 0.9, beta = 1.0 - 0.9;
    float best = 0.f;
    int idx = 0;
#pragma unroll
    for (int i = 0; i < 10; i++) {
      const double o = (double)(float)sm[i];
      const float f = (float)__dadd_rn(__dmul_rn(alpha, (double)pf[i]), __dmul_rn(beta, o));
      filt[clip * 10 + i] = f;
      if (i == 0 || f > best) { best = f; idx = i; }
    }
    if (detect) detect[clip] = best > 0.5f ? idx : -1;
  }
};

struct Job {
  const uint8_t *image;
  const int8_t *net_in;
  long long n, nb;      // clips, batches
  int row0;             // ring offset of the input rows (0 for plain batches)
  int nbc;              // batches of this CTA: blockIdx.x + k * gridDim.x, k < nbc
  bool aligned;         // net_in is 16-byte aligned: full batches come in by bulk copy
  int row_begin;        // first input row the conv1 units of this kernel read (kIncrFirstRow in the incremental kernel: only C1(2, x) runs)
  __device__ __forceinline__ long long batch(int k) const { return (long long)blockIdx.x + (long long)k * gridDim.x; }
  __device__ __forceinline__ bool bulk(int k) const { return aligned && (batch(k) + 1) * NC <= n; }
};

// input layout step for this CTA's k-th batch: raw [clip][403] -> c0[row & 1][row >> 1][clip][16]; tid in 0..nthreads-1
__device__ __forceinline__ void transpose_in(uint8_t *smem, const Job &J, int k, uint32_t raw_bar, int tid, int nthreads = 128) {
  const long long clip0 = J.batch(k) * NC;
  const int nclips = (int)min((long long)NC, J.n - clip0);
  const bool bulk = J.bulk(k);
  if (bulk) mbar_wait(raw_bar, (uint32_t)k & 1u);
  const int clip = tid & 31;
#pragma unroll 3
  for (int row = J.row_begin + (tid >> 5); row < kInRowsUsed; row += nthreads >> 5) {
    {
      const uint32_t a = raw_offset(clip, row, J.row0);
      uint32_t w[4];
      Vec16 v;
      if (bulk) {
        const uint32_t *wp = reinterpret_cast<const uint32_t *>(smem + OFF_RAW) + (a >> 2);
#pragma unroll
        for (int j = 0; j < 4; j++) w[j] = wp[j];
        v = input_chunk(w, a);
      } else {   // partial last batch or unaligned caller buffer: byte loads straight from HBM
        w[0] = w[1] = w[2] = w[3] = 0;
        if (clip < nclips) {
          const uint8_t *src = reinterpret_cast<const uint8_t *>(J.net_in) + clip0 * kNetIn + raw_offset(clip, row, J.row0);
#pragma unroll
          for (int j = 0; j < 13; j++) w[j >> 2] |= (uint32_t)src[j] << (8 * (j & 3));
        }
        v = Vec16{w[0], w[1], w[2], w[3]};
      }
      *reinterpret_cast<uint4 *>(smem + c0_offset(row, clip)) = make_uint4(v.x, v.y, v.z, v.w);
    }
  }
}

// incremental kernel: the two older rows of a layer for one batch, from the row cache into their places in the activation buffers, as
// bulk copies onto one mbarrier (issued by one thread).  conv3: rows 0, 1 (4 planes x 5 px); pooled conv2: rows 2, 3 (2 planes x 7 px)
__device__ __forceinline__ void start_cached_c3(uint32_t sbase, const uint8_t *cb, int set, int older, int newer, uint32_t bar) {
  mbar_expect_tx(bar, CACHE_C3_LOAD);
#pragma unroll
  for (int row = 0; row < 2; row++)
#pragma unroll
    for (int plane = 0; plane < 4; plane++)
      bulk_g2s(sbase + OFF_C3 + (uint32_t)plane * PLANE3 + (uint32_t)(row * 5) * PX,
               cb + cache_c3_off(set, row ? newer : older) + (uint32_t)plane * CACHE_C3_ROW, CACHE_C3_ROW, bar);
}
__device__ __forceinline__ void start_cached_c2(uint32_t sbase, const uint8_t *cb, int set, int older, int newer, uint32_t bar) {
  mbar_expect_tx(bar, CACHE_C2_LOAD);
#pragma unroll
  for (int row = 0; row < 2; row++)
#pragma unroll
    for (int plane = 0; plane < 2; plane++)
      bulk_g2s(sbase + OFF_C2P + (uint32_t)plane * PLANE2 + (uint32_t)((2 + row) * 7) * PX,
               cb + cache_c2_off(set, row ? newer : older) + (uint32_t)plane * CACHE_C2_ROW, CACHE_C2_ROW, bar);
}

__constant__ Unit c_sched[kUnits];
__constant__ Unit c_sched_i[kUnitsI];   // k2::kSchedI, the incremental kernel's   // k2::kSched for the epilogue loop, uploaded once by nnom_umma_configure (304 bytes: stays cached)

__device__ __forceinline__ uint64_t desc_from_lo(uint32_t lo) { return (uint64_t)lo | ((uint64_t)(8u | (1u << 14)) << 32); }

// The issue path is unrolled at compile time: every descriptor is (shared-memory base >> 4) + an immediate, the code is a
// straight line of ~7 instructions per MMA that ptxas software-pipelines.  (Working descriptors out at run time -- from tables
// in constant memory, in closed form per unit, or broadcast from registers -- made the single issuing warp the bottleneck:
// 150-200 cycles per MMA against ~45 of tensor time; profiles/r02b_k2_trace.txt.)
template <int POS, int M> struct IssueMmas {
  static __device__ __forceinline__ void go(uint32_t tm, uint32_t sb4) {
    constexpr Unit U = kSched[POS];
    if constexpr (M < mma_count(U)) {
      constexpr MmaOp op = mma_op(U, M);
      mma_i8(tm + op.col, desc_from_lo(desc_lo(op.a_off, op.a_lbo) + sb4), desc_from_lo(desc_lo(op.b_off, op.b_lbo) + sb4), idesc_i8(op.n), op.acc);
      IssueMmas<POS, M + 1>::go(tm, sb4);
    }
  }
};

template <int POS, bool TRACE> struct IssueUnits {
  static __device__ __forceinline__ void go(const Job &J, uint32_t bars, uint32_t bar_raw, uint32_t sbase, uint32_t sb4, uint32_t tmem, int it,
                                            uint32_t &gpos, int lane, long long *trace) {
    constexpr Unit U = kSched[POS];
    const uint32_t slot = gpos & (kSlots - 1), par = (gpos >> 2) & 1;
    if (TRACE && trace && it == 1 && lane == 0) trace[4 * POS + 0] = clock64();
    mbar_wait(bars + 32 + 8 * slot, par ^ 1);            // the team of this slot has drained its previous unit
    tc_fence_after();
    if (TRACE && trace && it == 1 && lane == 0) trace[4 * POS + 1] = clock64();
    // raw (batch `it`'s bytes) was consumed by the epilogue of unit kPosTranspose of the previous iteration, whose slot this wait covered
    if (POS == kPosNextRaw && it >= 1 && it + 1 < J.nbc && J.bulk(it + 1) && elect_one()) {
      mbar_expect_tx(bar_raw, NC * kNetIn);
      bulk_g2s(sbase + OFF_RAW, J.net_in + J.batch(it + 1) * NC * kNetIn, NC * kNetIn, bar_raw);
    }
    if (U.lag ? it > 0 : it < J.nbc) IssueMmas<POS, 0>::go(tmem + slot * kSlotCols, sb4);
    umma_commit(bars + 8 * slot);
    gpos++;
    if constexpr (POS + 1 < kUnits) IssueUnits<POS + 1, TRACE>::go(J, bars, bar_raw, sbase, sb4, tmem, it, gpos, lane, trace);
  }
};

// TRACE: block 0 records clock64() per unit of its second iteration -- issuing lane: [4 pos] = before the slot wait, [4 pos + 1] =
// slot free; thread 0: [4 pos + 2] = accumulators ready, [4 pos + 3] = slot released (tools/k2_trace.py prints the timeline)
template <bool TRACE>
__global__ void __launch_bounds__(kThreads, 1)
nnom_q7_umma_kernel(const uint8_t *image, const K2Launch K) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int8_t *net_in = K.net_in;
  const long long n = K.n;
  int8_t *logits = K.logits, *softmax = K.softmax;
  const int softmax_mode = K.softmax_mode;
  int *class_counts = K.class_counts;
  long long *trace = (TRACE && blockIdx.x == 0) ? K.trace : nullptr;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t sbase = smem_u32(smem), bars = sbase + OFF_BAR;
  // mbarriers: full[slot] = bars + 8 slot, empty[slot] = bars + 32 + 8 slot, raw = bars + 64, weights = bars + 72
  const uint32_t bar_raw = bars + 64, bar_wts = bars + 72;
  volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(smem + OFF_BAR + 80);
  int *hist = reinterpret_cast<int *>(smem + OFF_BAR + 96);
  Job J;
  J.image = image; J.net_in = net_in; J.n = n; J.nb = (n + NC - 1) / NC; J.row0 = K.row0;
  J.nbc = (long long)blockIdx.x < J.nb ? (int)((J.nb - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
  J.aligned = (reinterpret_cast<uintptr_t>(net_in) & 15) == 0;
  J.row_begin = 0;

  if (warp == kEpiWarps) {
    if (lane == 0) {
      for (int s = 0; s < kSlots; s++) {
        mbar_init(bars + 8 * s, 1);
        mbar_init(bars + 32 + 8 * s, 128);                      // one team of four warps per slot
      }
      mbar_init(bar_raw, 1); mbar_init(bar_wts, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      mbar_expect_tx(bar_wts, IMAGE_BYTES);
      bulk_g2s(sbase, image, IMAGE_BYTES, bar_wts);
      if (J.nbc > 0 && J.bulk(0)) {
        mbar_expect_tx(bar_raw, NC * kNetIn);
        bulk_g2s(sbase + OFF_RAW, net_in + J.batch(0) * NC * kNetIn, NC * kNetIn, bar_raw);
      }
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(bars + 80), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid < 10) hist[tid] = 0;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  mbar_wait(bar_wts, 0);                                       // weights and parameters are in shared memory
  if (warp < kEpiWarps && J.nbc > 0) transpose_in(smem, J, 0, bar_raw, tid, kEpiWarps * 32);
  fence_async_smem();
  __syncthreads();

  if (J.nbc > 0) {
    if (warp == kEpiWarps) {
      // ---- issuing warp (converged; elect.sync picks the lane that issues) ------------------------------------------------
      if (J.nbc > 1 && J.bulk(1) && elect_one()) {
        mbar_expect_tx(bar_raw, NC * kNetIn);
        bulk_g2s(sbase + OFF_RAW, net_in + J.batch(1) * NC * kNetIn, NC * kNetIn, bar_raw);
      }
      const uint32_t sb4 = sbase >> 4;
      uint32_t gpos = 0;
#pragma unroll 1
      for (int it = 0; it <= J.nbc; it++) IssueUnits<0, TRACE>::go(J, bars, bar_raw, sbase, sb4, tmem, it, gpos, lane, trace);
    } else {
      // ---- epilogue warps: team = warp >> 2 drains the units of TMEM slot `team`; within a team warp q owns lane quarter q ------
      DevCtx c;
      c.smem = smem; c.lane = lane; c.q = warp & 3; c.g = 0;
      c.clip0 = 0; c.nclips = 0; c.logits = logits; c.softmax = softmax; c.softmax_mode = softmax_mode; c.hist = hist;
      c.t_ld = 0; c.tracing = false; c.filt = K.filt; c.detect = K.detect; c.counting = K.class_counts != nullptr;
      c.cache_batch = nullptr; c.cache_w_c2 = kNoCache; c.cache_w_c3 = kNoCache;
      const uint32_t team = warp >> 2, ttid = tid & 127;
      const uint32_t tmem_q = tmem + ((uint32_t)(c.q * 32) << 16) + team * kSlotCols;
      const uint32_t bar_full = bars + 8 * team, bar_empty = bars + 32 + 8 * team;
      const uint32_t total = (uint32_t)(J.nbc + 1) * kUnits;
      c.taddr = tmem_q;
#pragma unroll 1
      for (uint32_t gpos = team; gpos < total; gpos += kSlots) {
        const int it = (int)(gpos / kUnits), pos = (int)(gpos - (uint32_t)it * kUnits);
        const Unit U = c_sched[pos];
        mbar_wait(bar_full, (gpos >> 2) & 1);                  // the unit's MMAs (and every earlier one) have completed
        tc_fence_after();
        const bool stamp = TRACE && trace && it == 1 && ttid == 0;
        if (stamp) trace[4 * pos + 2] = clock64();
        if (TRACE) c.tracing = stamp && pos == kPosTranspose;  // fine stamps of one unit: [4 kUnits] last ld landed, [+1] stored, [+2] fenced
        if (U.lag ? it > 0 : it < J.nbc) {
          if (U.kind == DN) {
            if (c.q == 0) {
              c.clip0 = J.batch(it - 1) * NC;
              c.nclips = (int)min((long long)NC, J.n - c.clip0);
              run_dense(c);
            }
          } else {
            epilogue_conv(c, U);
          }
        }
        if (stamp && c.tracing) { trace[4 * kUnits] = c.t_ld; trace[4 * kUnits + 1] = clock64(); }
        if (pos == kPosTranspose && it + 1 < J.nbc && c.q > 0) transpose_in(smem, J, it + 1, bar_raw, (int)ttid - 32, 96);   // warp 0 of the team runs the dense epilogue
        fence_async_smem();   // this thread's shared-memory stores -> visible to the MMAs that will read them
        tc_fence_before();
        if (stamp && c.tracing) trace[4 * kUnits + 2] = clock64();
        mbar_arrive(bar_empty);
        if (stamp) trace[4 * pos + 3] = clock64();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(512) : "memory");
  if (class_counts && tid < 10 && hist[tid]) atomicAdd(&class_counts[tid], hist[tid]);
}

// ---- incremental kernel (streaming, k2_umma_plan.h "incremental sliding-window mode") -------------------------------------------
// Same machinery as above on the eight-unit schedule kSchedI: one pipeline stage per iteration, five batches in flight.
template <int POS, int M> struct IssueMmasI {
  static __device__ __forceinline__ void go(uint32_t tm, uint32_t sb4) {
    constexpr Unit U = kSchedI[POS];
    if constexpr (M < mma_count(U)) {
      constexpr MmaOp op = mma_op(U, M);
      mma_i8(tm + op.col, desc_from_lo(desc_lo(op.a_off, op.a_lbo) + sb4), desc_from_lo(desc_lo(op.b_off, op.b_lbo) + sb4), idesc_i8(op.n), op.acc);
      IssueMmasI<POS, M + 1>::go(tm, sb4);
    }
  }
};

template <int POS, bool TRACE> struct IssueUnitsI {
  static __device__ __forceinline__ void go(const Job &J, uint32_t bars, uint32_t bar_raw, uint32_t sbase, uint32_t sb4, uint32_t tmem, int it, int lane,
                                            long long *trace) {
    constexpr Unit U = kSchedI[POS];
    constexpr uint32_t slot = slot_i(POS), par = POS >> 2;   // 8 units on 4 slots, two uses per slot and iteration: fixed per position
    const int b = it - U.lag;                               // this unit's batch (of this CTA)
    const bool active = b >= 0 && b < J.nbc;
    if (TRACE && trace && it == 5 && lane == 0) trace[4 * POS + 0] = clock64();
    mbar_wait(bars + 32 + 8 * slot, par ^ 1);
    tc_fence_after();
    if (TRACE && trace && it == 5 && lane == 0) trace[4 * POS + 1] = clock64();
    // the team of conv3 tile 3 (this slot's previous user) has laid out this batch's input rows: the raw buffer (batch `it`) has been consumed
    if (POS == kPosI_Conv1 && it >= 1 && it + 1 < J.nbc && J.bulk(it + 1) && elect_one()) {
      mbar_expect_tx(bar_raw, NC * kNetIn);
      bulk_g2s(sbase + OFF_RAW, J.net_in + J.batch(it + 1) * NC * kNetIn, NC * kNetIn, bar_raw);
    }
    if (POS == kPosI_Conv4 && active) mbar_wait(bars + BAR_CACHE, (uint32_t)b & 1u);    // conv3's cached rows of this batch have landed
    if (POS == kPosI_Conv3 && active) mbar_wait(bars + BAR_CACHE2, (uint32_t)b & 1u);   // conv2's cached rows of this batch have landed
    if (active) IssueMmasI<POS, 0>::go(tmem + slot * kSlotCols, sb4);
    umma_commit(bars + 8 * slot);
    if constexpr (POS + 1 < kUnitsI) IssueUnitsI<POS + 1, TRACE>::go(J, bars, bar_raw, sbase, sb4, tmem, it, lane, trace);
  }
};

template <bool TRACE>
__global__ void __launch_bounds__(kThreads, 1)
nnom_q7_umma_incr_kernel(const uint8_t *image, const K2Launch K) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int8_t *net_in = K.net_in;
  const long long n = K.n;
  long long *trace = (TRACE && blockIdx.x == 0) ? K.trace : nullptr;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t sbase = smem_u32(smem), bars = sbase + OFF_BAR;
  const uint32_t bar_raw = bars + 64, bar_wts = bars + 72;
  volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(smem + OFF_BAR + 80);
  int *hist = reinterpret_cast<int *>(smem + OFF_BAR + 96);
  Job J;
  J.image = image; J.net_in = net_in; J.n = n; J.nb = (n + NC - 1) / NC; J.row0 = K.row0;
  J.nbc = (long long)blockIdx.x < J.nb ? (int)((J.nb - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
  J.aligned = (reinterpret_cast<uintptr_t>(net_in) & 15) == 0;
  J.row_begin = kIncrFirstRow;

  if (warp == kEpiWarps) {
    if (lane == 0) {
      for (int s = 0; s < kSlots; s++) {
        mbar_init(bars + 8 * s, 1);
        mbar_init(bars + 32 + 8 * s, 128);
      }
      mbar_init(bar_raw, 1); mbar_init(bar_wts, 1); mbar_init(bars + BAR_CACHE, 1); mbar_init(bars + BAR_CACHE2, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      mbar_expect_tx(bar_wts, IMAGE_BYTES);
      bulk_g2s(sbase, image, IMAGE_BYTES, bar_wts);
      if (J.nbc > 0 && J.bulk(0)) {
        mbar_expect_tx(bar_raw, NC * kNetIn);
        bulk_g2s(sbase + OFF_RAW, net_in + J.batch(0) * NC * kNetIn, NC * kNetIn, bar_raw);
      }
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(bars + 80), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid < 10) hist[tid] = 0;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  mbar_wait(bar_wts, 0);
  if (warp < kEpiWarps && J.nbc > 0) transpose_in(smem, J, 0, bar_raw, tid, kEpiWarps * 32);
  fence_async_smem();
  __syncthreads();

  if (J.nbc > 0) {
    const int n_it = J.nbc + kMaxLagI;
    if (warp == kEpiWarps) {
      if (J.nbc > 1 && J.bulk(1) && elect_one()) {
        mbar_expect_tx(bar_raw, NC * kNetIn);
        bulk_g2s(sbase + OFF_RAW, net_in + J.batch(1) * NC * kNetIn, NC * kNetIn, bar_raw);
      }
      const uint32_t sb4 = sbase >> 4;
#pragma unroll 1
      for (int it = 0; it < n_it; it++) IssueUnitsI<0, TRACE>::go(J, bars, bar_raw, sbase, sb4, tmem, it, lane, trace);
    } else {
      DevCtx c;
      c.smem = smem; c.lane = lane; c.q = warp & 3; c.g = 0;
      c.clip0 = 0; c.nclips = 0; c.logits = K.logits; c.softmax = K.softmax; c.softmax_mode = K.softmax_mode; c.hist = hist;
      c.t_ld = 0; c.tracing = false; c.filt = K.filt; c.detect = K.detect; c.counting = K.class_counts != nullptr;
      c.cache_batch = nullptr;
      c.cache_w_c2 = cache_c2_off(K.cache_set, K.cache_write);
      c.cache_w_c3 = cache_c3_off(K.cache_set, K.cache_write);
      const uint32_t team = warp >> 2, ttid = tid & 127;
      c.taddr = tmem + ((uint32_t)(c.q * 32) << 16) + team * kSlotCols;
      const uint32_t bar_full = bars + 8 * team, bar_empty = bars + 32 + 8 * team;
      const uint32_t total = (uint32_t)n_it * kUnitsI;
      (void)total;
#pragma unroll 1
      for (uint32_t j = 0; j < 2u * (uint32_t)n_it; j++) {     // the j-th use of this team's slot
        const int it = (int)(j >> 1), pos = team_pos_i((int)team, (int)(j & 1));
        const Unit U = c_sched_i[pos];
        const int b = it - U.lag;
        const bool active = b >= 0 && b < J.nbc;
        mbar_wait(bar_full, j & 1);
        tc_fence_after();
        const bool stamp = TRACE && trace && it == 5 && ttid == 0;
        if (stamp) trace[4 * pos + 2] = clock64();
        if (TRACE) c.tracing = stamp && pos == kPosI_Dense;      // fine stamps of the dense unit: [4 kUnitsI] accumulators in registers, [+1] run_dense done
        if (active) {
          c.cache_batch = K.cache + (size_t)J.batch(b) * CACHE_BATCH_BYTES;
          if (U.kind == DN) {
            if (c.q == 0) {
              c.clip0 = J.batch(b) * NC;
              c.nclips = (int)min((long long)NC, J.n - c.clip0);
              run_dense(c);
            }
          } else {
            epilogue_conv(c, U);
          }
          if (stamp && c.tracing) { trace[4 * kUnitsI] = c.t_ld; trace[4 * kUnitsI + 1] = clock64(); }
          // every MMA issued before this unit has completed, the readers of the previous batch's cached rows among them: fetch this batch's
          if (pos == kPosI_Conv3 && ttid == 32) start_cached_c3(sbase, c.cache_batch, K.cache_set, K.cache_older, K.cache_newer, bars + BAR_CACHE);
          if (pos == kPosI_Conv2 && ttid == 32) start_cached_c2(sbase, c.cache_batch, K.cache_set, K.cache_older, K.cache_newer, bars + BAR_CACHE2);
        }
        // the two warps of conv3 tile 3's team whose positions are wrap-around pixels (no jobs) lay out batch `it`'s input rows (batch 0 was laid
        // out before the loop): conv1 of batch it - 1 is done with c0, and the first conv1 unit of this iteration is the slot's next user
        if (pos == kPosI_Conv3 && it >= 1 && it < J.nbc && c.q < 2) transpose_in(smem, J, it, bar_raw, (int)ttid, 64);
        fence_async_smem();
        tc_fence_before();
        mbar_arrive(bar_empty);
        if (stamp) trace[4 * pos + 3] = clock64();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(512) : "memory");
  if (K.class_counts && tid < 10 && hist[tid]) atomicAdd(&K.class_counts[tid], hist[tid]);
}

}  // namespace

// ---- host side -----------------------------------------------------------------------------------------------------------
cudaError_t nnom_umma_configure(void) {
  cudaError_t e = cudaMemcpyToSymbol(c_sched, kSched, sizeof(kSched));
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_sched_i, kSchedI, sizeof(kSchedI));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(nnom_q7_umma_incr_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(nnom_q7_umma_incr_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(nnom_q7_umma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(nnom_q7_umma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
}

// builds the weight image of k2_umma_plan.h from the five layers and uploads it; *out_image is a device pointer
int nnom_umma_upload(const int8_t *const w[5], const int8_t *const b[5], const int rs[5], const int ls[5], uint8_t **out_image) {
  uint8_t *host = (uint8_t *)malloc(IMAGE_BYTES);
  if (!host) return EDISON_E_NOMEM;
  build_image(host, w, b, rs, ls);
  uint8_t *d = nullptr;
  cudaError_t e = cudaMalloc(&d, IMAGE_BYTES);
  if (e == cudaSuccess) e = cudaMemcpy(d, host, IMAGE_BYTES, cudaMemcpyHostToDevice);
  free(host);
  if (e != cudaSuccess) {
    if (d) cudaFree(d);
    return EDISON_E_CUDA;
  }
  *out_image = d;
  return EDISON_OK;
}

size_t nnom_umma_cache_bytes(long long n) { return (size_t)((n + NC - 1) / NC) * CACHE_BATCH_BYTES; }


cudaError_t launch_nnom_q7_umma(const uint8_t *image, const K2Launch &k, int sm_count, cudaStream_t stream) {
  if (k.n <= 0) return cudaSuccess;
  const long long nb = (k.n + NC - 1) / NC;
  const unsigned ctas = (unsigned)(nb < (long long)sm_count ? nb : (long long)sm_count);
  if (k.cache) {   // incremental sliding-window mode
    if (k.trace) nnom_q7_umma_incr_kernel<true><<<ctas, kThreads, SMEM_BYTES, stream>>>(image, k);
    else nnom_q7_umma_incr_kernel<false><<<ctas, kThreads, SMEM_BYTES, stream>>>(image, k);
  } else if (k.trace) nnom_q7_umma_kernel<true><<<ctas, kThreads, SMEM_BYTES, stream>>>(image, k);
  else nnom_q7_umma_kernel<false><<<ctas, kThreads, SMEM_BYTES, stream>>>(image, k);
  count_launch();
  return cudaGetLastError();
}

}  // namespace edison
