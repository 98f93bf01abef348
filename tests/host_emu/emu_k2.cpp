// emu_k2.cpp -- host emulation of the K2 tcgen05 plan (TEST INFRASTRUCTURE).
//
// Runs edison_b200/csrc/k2_umma_plan.h -- the same MMA list, the same epilogue templates, the same weight image and the same
// unit order as the CUDA kernel (nnom_q7_umma.cu) -- on a byte array standing in for shared memory and an int32 array
// standing in for TMEM.  An MMA is emulated with the descriptor addressing that tools/ubench/umma_probe.cu verified on a
// B200 (A[m][k] = smem[start + (m / 8) * 128 + (m % 8) * 16 + (k / 16) * LBO + k % 16]).  `eager` moves every unit's MMAs in
// front of the epilogues of the three units before it, the other extreme of what the slot ring allows, so that a
// write-after-read hazard between an epilogue and MMAs already in flight would show as a mismatch in one of the two orders.
// It exists to check index arithmetic against the oracle in a container without a GPU; it is not part of
// libedison_b200.so and never produces results: the -m gpu parity tests run the real kernel.
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "k2_umma_plan.h"
#include "nnom_blob.h"

using namespace edison;
using namespace edison::k2;

namespace {

struct Emu {
  std::vector<uint8_t> smem;
  std::vector<int32_t> tmem;   // [128 lanes][512 columns]
  Emu() : smem(SMEM_BYTES + 4096, 0xA5), tmem(128 * 512, 0x5a5a5a5a) {}

  // one MMA, decoded from what the kernel hands to tcgen05.mma (descriptor low words, instruction descriptor)
  void mma(const MmaOp &op, int slot) {
    const uint32_t a_lo = desc_lo(op.a_off, op.a_lbo), b_lo = desc_lo(op.b_off, op.b_lbo), idesc = idesc_i8(op.n);
    const uint32_t a_off = (a_lo & 0x3fff) << 4, a_lbo = ((a_lo >> 16) & 0x3fff) << 4;
    const uint32_t b_off = (b_lo & 0x3fff) << 4, b_lbo = ((b_lo >> 16) & 0x3fff) << 4;
    const uint32_t n = ((idesc >> 17) & 0x3f) << 3, col = op.col, acc = op.acc;
    for (int m = 0; m < 128; m++)
      for (uint32_t j = 0; j < n; j++) {
        int32_t &d = tmem[m * 512 + slot * kSlotCols + col + j];
        int32_t s = acc ? d : 0;
        for (int k = 0; k < 32; k++) {
          const uint32_t aa = a_off + (m / 8) * 128 + (m % 8) * 16 + (k / 16) * a_lbo + (k % 16);
          const uint32_t ba = b_off + (j / 8) * 128 + (j % 8) * 16 + (k / 16) * b_lbo + (k % 16);
          s += (int)(int8_t)smem[aa] * (int)(int8_t)smem[ba];
        }
        d = s;
      }
  }
};

struct Ctx {
  Emu *e;
  int slot, lane, q, g;
  long long clip0;
  int nclips;
  int8_t *logits, *softmax;
  int softmax_mode;
  long long *hist;
  float *filt = nullptr;
  bool counting = true;
  uint32_t cache_w_c2 = kNoCache, cache_w_c3 = kNoCache;
  uint8_t *cache_batch = nullptr;      // incremental mode: this unit's batch in the row cache
  void st16_cache(uint32_t off, Vec16 v) { memcpy(cache_batch + off, &v, 16); }
  void prefetch(long long) {}
  void post(long long, const int8_t *) {}
  void ld16(uint32_t col, uint32_t (&v)[16]) {
    for (int j = 0; j < 16; j++) v[j] = (uint32_t)e->tmem[(q * 32 + lane) * 512 + slot * kSlotCols + col + j];
  }
  void ld_wait() {}
  void st16(uint32_t off, Vec16 v) { memcpy(&e->smem[off], &v, 16); }
  int par(int i) { int32_t v; memcpy(&v, &e->smem[OFF_PAR + 4 * i], 4); return v; }
  Int4 par4(int i) { Int4 v; memcpy(&v, &e->smem[OFF_PAR + 4 * i], 16); return v; }
  void count(int cls) { if (hist) hist[cls]++; }
};

void transpose_in(Emu &E, const int8_t *net_in, long long n, long long batch, int row_begin = 0) {
  const long long clip0 = batch * NC;
  const int nclips = (int)((n - clip0) < NC ? (n - clip0) : NC);
  memset(&E.smem[OFF_RAW], 0x77, RAW_BYTES);
  memcpy(&E.smem[OFF_RAW], net_in + clip0 * kNetIn, (size_t)nclips * kNetIn);
  for (int row = row_begin; row < kInRowsUsed; row++)
    for (int clip = 0; clip < NC; clip++) {
      const uint32_t a = raw_offset(clip, row, 0);
      uint32_t w[4];
      memcpy(w, &E.smem[OFF_RAW + (a & ~3u)], 16);
      const Vec16 v = input_chunk(w, a);
      memcpy(&E.smem[c0_offset(row, clip)], &v, 16);
    }
}

void run_mma(Emu &E, int pos, long long gpos, bool head, bool tail) {
  const Unit U = kSched[pos];
  if (!(U.lag ? tail : head)) return;
  for (int m = 0; m < mma_count(U); m++) E.mma(mma_op(U, m), (int)(gpos & (kSlots - 1)));
}

void run_epi(Emu &E, int pos, long long gpos, bool head, bool tail, const int8_t *net_in, long long n, long long batch_of_it, long long next_batch,
             int8_t *logits, int8_t *softmax, int mode, long long *hist) {
  const Unit U = kSched[pos];
  if (U.lag ? tail : head) {
    for (int q = 0; q < 4; q++)   // the four warps of the slot's team
      for (int lane = 0; lane < 32; lane++) {
        Ctx c;
        c.e = &E; c.slot = (int)(gpos & (kSlots - 1)); c.lane = lane; c.q = q; c.g = 0;
        const long long b = U.lag ? batch_of_it - 1 : batch_of_it;   // single-CTA emulation: consecutive batches
        c.clip0 = b * NC;
        c.nclips = (int)((n - c.clip0) < NC ? (n - c.clip0) : NC);
        c.logits = logits; c.softmax = softmax; c.softmax_mode = mode; c.hist = hist;
        if (U.kind == DN) {
          if (q == 0) run_dense(c);
        } else {
          epilogue_conv(c, U);
        }
      }
  }
  if (pos == kPosTranspose && next_batch >= 0) transpose_in(E, net_in, n, next_batch);
}

}  // namespace

// net_in [n][403] -> logits [n][10], softmax [n][10], hist [10] (any may be null); returns 0 or a negative code
extern "C" int emu_k2_run(const char *blob_path, const int8_t *net_in, long long n, int8_t *logits, int8_t *softmax, int mode, long long *hist,
                          int eager) {
  KwsBlob blob;
  char err[256] = "";
  int r = read_kws_blob(blob_path, blob, err, sizeof err);
  if (r) { fprintf(stderr, "emu_k2_run: %s\n", err); return r; }
  const int8_t *w[5], *b[5];
  int rs[5], ls[5];
  for (int i = 0; i < 5; i++) { w[i] = blob.L[i].w; b[i] = blob.L[i].b; rs[i] = (int)blob.L[i].h[5]; ls[i] = (int)blob.L[i].h[6]; }
  Emu E;
  build_image(E.smem.data(), w, b, rs, ls);
  if (hist) memset(hist, 0, 10 * sizeof(long long));
  int total_mma = 0;
  for (int p = 0; p < kUnits; p++) total_mma += mma_count(kSched[p]);
  if (total_mma != kMmaPerIter) { fprintf(stderr, "emu_k2_run: kMmaPerIter is %d, the schedule has %d\n", kMmaPerIter, total_mma); return -1; }
  const long long nb = (n + NC - 1) / NC;
  if (nb == 0) return 0;
  transpose_in(E, net_in, n, 0);
  // lazy: MMA(u) right before epilogue(u).  eager: the issuing thread runs as far ahead as the slot ring lets it, i.e. the MMAs of
  // unit u + 3 execute before the epilogue of unit u (they only had to wait for the epilogue of u - 1).
  const long long total = (nb + 1) * kUnits;
  const int ahead = eager ? kSlots - 1 : 0;
  long long issued = 0;
  for (long long u = 0; u < total; u++) {
    for (; issued <= u + ahead && issued < total; issued++) {
      const long long it = issued / kUnits;
      run_mma(E, (int)(issued % kUnits), issued, it < nb, it > 0);
    }
    const long long it = u / kUnits;
    const bool head = it < nb, tail = it > 0;
    run_epi(E, (int)(u % kUnits), u, head, tail, net_in, n, it, head && it + 1 < nb ? it + 1 : -1, logits, softmax, mode, hist);
  }
  return 0;
}

extern "C" unsigned emu_k2_image_bytes(void) { return IMAGE_BYTES; }
extern "C" unsigned emu_k2_smem_bytes(void) { return SMEM_BYTES; }

// ---- incremental mode (k2_umma_plan.h kSchedI; nnom_q7_umma_incr_kernel) --------------------------------------------------------------
// One CTA, consecutive batches, one hop.  The row cache [nb][CACHE_BATCH_BYTES] persists across calls (hops).  Order of execution:
//   lazy : every unit's MMAs right before its epilogue;
//   eager: before each epilogue, the issuing thread runs as far ahead as the kernel lets it -- a unit's MMAs may run once the previous
//          user of its TMEM slot (slot_i) has been drained and, for the units that read cached rows, once the copy started by an earlier
//          epilogue has been made -- the other extreme, so that a missing dependency shows as a mismatch in one of the two orders.
namespace {

struct IncrRun {
  Emu &E;
  const int8_t *net_in;
  long long n, nb;
  uint8_t *cache;
  int set, older, newer, write;
  int8_t *softmax;
  int mode;
  long long n_units;
  std::vector<char> mma_done, drained, c2_copied, c3_copied;   // per unit number / per batch

  const Unit &unit(long long u) const { return kSchedI[u % kUnitsI]; }
  long long batch_of(long long u) const { return u / kUnitsI - unit(u).lag; }
  bool active(long long u) const { const long long b = batch_of(u); return b >= 0 && b < nb; }
  // the previous unit that used the same TMEM slot
  long long prev_same_slot(long long u) const {
    for (long long v = u - 1; v >= 0; v--)
      if (slot_i((int)(v % kUnitsI)) == slot_i((int)(u % kUnitsI))) return v;
    return -1;
  }
  bool may_issue(long long u) const {
    const long long p = prev_same_slot(u);
    if (p >= 0 && !drained[p]) return false;
    const int pos = (int)(u % kUnitsI);
    if (active(u) && pos == kPosI_Conv4 && !c3_copied[batch_of(u)]) return false;
    if (active(u) && pos == kPosI_Conv3 && !c2_copied[batch_of(u)]) return false;
    return true;
  }
  void do_mma(long long u) {
    if (active(u))
      for (int m = 0; m < mma_count(unit(u)); m++) E.mma(mma_op(unit(u), m), slot_i((int)(u % kUnitsI)));
    mma_done[u] = 1;
  }
  void copy_rows(uint32_t dst, const uint8_t *src, uint32_t bytes) { memcpy(&E.smem[dst], src, bytes); }
  void do_epilogue(long long u) {
    const int pos = (int)(u % kUnitsI);
    const long long it = u / kUnitsI;
    const Unit U = unit(u);
    if (active(u)) {
      const long long b = batch_of(u);
      uint8_t *cb = cache + (size_t)b * CACHE_BATCH_BYTES;
      for (int q = 0; q < 4; q++)
        for (int lane = 0; lane < 32; lane++) {
          Ctx c;
          c.e = &E; c.slot = slot_i(pos); c.lane = lane; c.q = q; c.g = 0;
          c.clip0 = b * NC;
          c.nclips = (int)((n - c.clip0) < NC ? (n - c.clip0) : NC);
          c.logits = nullptr; c.softmax = softmax; c.softmax_mode = mode; c.hist = nullptr; c.counting = false;
          c.cache_batch = cb; c.cache_w_c2 = cache_c2_off(set, write); c.cache_w_c3 = cache_c3_off(set, write);
          if (U.kind == DN) {
            if (q == 0) run_dense(c);
          } else {
            epilogue_conv(c, U);
          }
        }
      // the bulk copies the kernel's teams start at this point (they have landed before the consumers are issued: c*_copied)
      if (pos == kPosI_Conv3) {
        for (int row = 0; row < 2; row++)
          for (int plane = 0; plane < 4; plane++)
            copy_rows(OFF_C3 + plane * PLANE3 + row * 5 * PX, cb + cache_c3_off(set, row ? newer : older) + plane * CACHE_C3_ROW, CACHE_C3_ROW);
        c3_copied[b] = 1;
      }
      if (pos == kPosI_Conv2) {
        for (int row = 0; row < 2; row++)
          for (int plane = 0; plane < 2; plane++)
            copy_rows(OFF_C2P + plane * PLANE2 + (2 + row) * 7 * PX, cb + cache_c2_off(set, row ? newer : older) + plane * CACHE_C2_ROW, CACHE_C2_ROW);
        c2_copied[b] = 1;
      }
    }
    if (pos == kPosI_Conv3 && it >= 1 && it < nb) transpose_in(E, net_in, n, it, kIncrFirstRow);
    drained[u] = 1;
  }
};

}  // namespace

// one hop of the incremental mode over n windows [n][403] (logical row order); cache: [ceil(n / 32)][CACHE_BATCH_BYTES], kept by the caller
// between hops; set / older / newer / write as stream.cu derives them from the hop number.  softmax [n][10].
extern "C" int emu_k2_incr_hop(const char *blob_path, const int8_t *net_in, long long n, uint8_t *cache, int set, int older, int newer, int write,
                               int8_t *softmax, int mode, int eager) {
  KwsBlob blob;
  char err[256] = "";
  int r = read_kws_blob(blob_path, blob, err, sizeof err);
  if (r) { fprintf(stderr, "emu_k2_incr_hop: %s\n", err); return r; }
  const int8_t *w[5], *b[5];
  int rs[5], ls[5];
  for (int i = 0; i < 5; i++) { w[i] = blob.L[i].w; b[i] = blob.L[i].b; rs[i] = (int)blob.L[i].h[5]; ls[i] = (int)blob.L[i].h[6]; }
  Emu E;
  build_image(E.smem.data(), w, b, rs, ls);
  const long long nb = (n + NC - 1) / NC;
  if (nb == 0) return 0;
  IncrRun R{E, net_in, n, nb, cache, set, older, newer, write, softmax, mode, (nb + kMaxLagI) * kUnitsI, {}, {}, {}, {}};
  R.mma_done.assign(R.n_units, 0); R.drained.assign(R.n_units, 0);
  R.c2_copied.assign(nb, 0); R.c3_copied.assign(nb, 0);
  transpose_in(E, net_in, n, 0, kIncrFirstRow);
  long long issued = 0;
  for (long long u = 0; u < R.n_units; u++) {
    if (eager)
      while (issued < R.n_units && R.may_issue(issued)) R.do_mma(issued++);
    if (!R.mma_done[u]) {
      if (!R.may_issue(u)) { fprintf(stderr, "emu_k2_incr_hop: unit %lld can never be issued (dependency cycle)\n", u); return -2; }
      while (issued <= u) R.do_mma(issued++);
    }
    R.do_epilogue(u);
  }
  return 0;
}

extern "C" unsigned emu_k2_cache_batch_bytes(void) { return CACHE_BATCH_BYTES; }
