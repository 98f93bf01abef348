// k2_umma_plan.h -- the static plan of K2 (int8 NNoM keyword-spotting net, firmware/src/ai/nnom/kws_nnom/weights.h:138-161)
// on the Blackwell tensor path: which tcgen05.mma instructions run, on which shared-memory bytes, in which order, and how
// each accumulator tile is turned into the next layer's activations.  Everything here is plain constexpr / inline C++ so
// that the SAME plan drives the CUDA kernel (nnom_q7_umma.cu) and a host emulation of it (tests/host_emu/emu_k2.cpp, which
// checks the index arithmetic against the oracle without a GPU).
//
// Layout idea ("clip-minor activations, convolution by descriptor strides"):
//   * a CTA owns a batch of NC = 32 clips; an activation tensor [H][W][C] of the batch lives in shared memory as
//       [C/16 planes][H*W pixels][32 clips][16 channels]            (512 B per pixel and plane)
//     so 8 consecutive 16-byte units are 8 clips at one pixel = one core matrix of a NO-SWIZZLE K-major UMMA operand, and
//     128 consecutive units (= the M = 128 rows of one MMA) are 4 consecutive pixels x 32 clips: the A operand of filter
//     tap (ky, kx) is the SAME array read from a start address shifted by (ky * W + kx) pixels -- no im2col, no copies;
//   * K = 32 int8 per instruction = two 16-byte chunks LBO bytes apart: two taps (C = 16), the two channel halves of one tap
//     (C = 32), or two of the four channel quarters (C = 64);
//   * outputs at the wrap-around pixels of an image row (x >= W_out) are computed and dropped;
//   * conv1 (one input channel) is a banded (Toeplitz) product instead: an A row is (clip, image row) with the 13 input
//     bytes of that row (padded to 16) as one chunk, N = (x_out, channel) = 9 x 16, B = the 5x5 kernel laid out along x;
//     even and odd conv rows are two column ranges of ONE MMA (same A rows, N = 2 x 48) and land in the same lane, so the
//     2x1 max pool is a max in registers;
//   * accumulators live in TMEM (lane = 32 * pixel-in-tile + clip, column = output channel), four slots of 128 columns;
//     16 epilogue warps read them back with tcgen05.ld (lane = clip), add the bias seed, shift, saturate, ReLU, pool,
//     pack 16 channels into one 16-byte store in the next layer's layout;
//   * every MMA descriptor and epilogue job is a closed form of (unit, index): the issue path is unrolled at compile time
//     (descriptors become immediates), the epilogue is one small loop.
// int32 accumulation is exact and order independent, so the result has the reference's bits
// (arm_convolve_HWC_q7_fast_nonsquare.c:74-375, arm_fully_connected_q7_opt.c:129-475, nnom_local.c:117-160, 425-484).
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define K2_HD __host__ __device__ __forceinline__
#else
#define K2_HD inline
#endif

namespace edison {
namespace k2 {

constexpr int NC = 32;                 // clips per batch
constexpr int kEpiWarps = 16;          // warps 0..15: epilogue (lane quarter = warp & 3, team = TMEM slot = warp >> 2); warp 16 issues
constexpr int kThreads = (kEpiWarps + 1) * 32;
constexpr uint32_t PX = NC * 16;       // bytes of one pixel in one 16-channel plane
constexpr int kNetIn = 403;

// ---- shared-memory map (byte offsets from the dynamic shared memory base) --------------------------------------------
// weight image (built once on the host, copied in with one bulk copy)
constexpr uint32_t W1_CHUNK = 288 * 16;                  // one K-chunk of a conv1 B matrix: rows [x third][tile E/O][px][c] x 16 bytes
constexpr uint32_t OFF_W1 = 0;                           // 3 chunk pairs j x 2 chunks
constexpr uint32_t OFF_W2 = OFF_W1 + 6 * W1_CHUNK;       // 13 blocks of [32][16]: Z k0w0 k0w1 k0w2 Z k1w0 k1w1 k1w2 Z k2w0 k2w1 k2w2 Z (k = kx, w = ky)
constexpr uint32_t W2_BLOCK = 512;
constexpr uint32_t OFF_W3 = OFF_W2 + 13 * W2_BLOCK;      // 9 x [2][64][16]
constexpr uint32_t OFF_W4 = OFF_W3 + 9 * 2048;           // 18 x [2][32][16]
constexpr uint32_t OFF_WD = OFF_W4 + 18 * 1024;          // 3 x [2][16][16]
constexpr uint32_t OFF_PAR = OFF_WD + 3 * 512;           // int32: seed1[16] seed2[32] seed3[64] seed4[32] seedd[16] rs[5] hi[5]
constexpr int PAR_SEED1 = 0, PAR_SEED2 = 16, PAR_SEED3 = 48, PAR_SEED4 = 112, PAR_SEEDD = 144, PAR_RS = 160, PAR_HI = 165;
constexpr uint32_t PAR_BYTES = 704;
constexpr uint32_t IMAGE_BYTES = OFF_PAR + PAR_BYTES;
// activations
constexpr uint32_t OFF_C0 = IMAGE_BYTES;                 // input rows: [parity][row / 2 (14)][clip][16]; rows 28..30 never matter
constexpr uint32_t C0_HALF = 14 * PX;
constexpr uint32_t OFF_RAW = OFF_C0 + 2 * C0_HALF;       // the batch's [32][403] input bytes as they lie in HBM (bulk copy target)
constexpr uint32_t RAW_BYTES = NC * kNetIn + 16;
constexpr uint32_t OFF_C1P = OFF_RAW + RAW_BYTES;        // pooled conv1: 12 rows x 9 px (+2 px read by zero-weight / dropped taps)
constexpr uint32_t OFF_C2P = OFF_C1P + 110 * PX;         // pooled conv2: 2 planes x (5 x 7 px + 1)
constexpr uint32_t PLANE2 = 36 * PX;
constexpr uint32_t OFF_C4 = OFF_C2P + 2 * PLANE2;        // conv4: 2 planes x 3 px; the dense tile's unused rows (pixels 3..5) read on into the
constexpr uint32_t PLANE4 = 3 * PX;                      //   next plane / conv3's buffer: they only feed accumulator rows nobody reads
constexpr uint32_t OFF_C3 = OFF_C4 + 2 * PLANE4;         // conv3: 4 planes x (3 x 5 px + 1)
constexpr uint32_t PLANE3 = 16 * PX;
constexpr uint32_t OFF_BAR = OFF_C3 + 4 * PLANE3;        // mbarriers full[4] empty[4] raw wts (80 B), TMEM base (+80), class histogram (+96, 40 B)
constexpr uint32_t SMEM_BYTES = OFF_BAR + 160;
static_assert(IMAGE_BYTES % 16 == 0 && RAW_BYTES % 16 == 0 && OFF_BAR % 16 == 0, "16-byte granularity");
static_assert(SMEM_BYTES <= 232448, "more than the 227 KB a CTA can have");

// ---- the schedule --------------------------------------------------------------------------------------------------------
// TMEM: four slots of 128 columns, unit number u (counted over the whole kernel) uses slot u & 3, and each slot has its own
// team of four epilogue warps (one per lane quarter), so up to four units are drained concurrently.  One loop iteration runs
// 21 units (3 or 4 epilogue jobs per warp each, so that no slot is held much longer than the others): the head of batch b (conv1, conv2) interleaved with the tail of batch b - 1 (conv3, conv4, dense: lag = 1), ordered
// so that every unit's inputs were produced at least four units earlier.  That distance is what makes the only wait of the
// issuing thread -- "my slot's previous unit u - 4 has been drained" -- sufficient: a producer d <= u - 5 was drained before
// unit d + 4 < u could be issued, and d = u - 4 is the slot wait itself.  An epilogue only starts after its own unit's MMAs
// (hence all earlier MMAs: the tensor pipe is in order) have completed, so it never overwrites data an older MMA still reads.
enum Kind { C1 = 0, C2 = 1, C3 = 2, C4 = 3, DN = 4 };
struct Unit { int kind, a, b, lag; };   // C1: a = t (conv rows 8t..8t+7), b = x third; C2: a = pooled row; C3: a = first tile, b = tiles
constexpr int kUnits = 21, kSlots = 4, kSlotCols = 128;
constexpr Unit kSched[kUnits] = {
    {C1, 0, 0, 0}, {C3, 0, 1, 1}, {C1, 0, 1, 0}, {C3, 1, 1, 1}, {C1, 0, 2, 0}, {C3, 2, 1, 1}, {C1, 1, 0, 0}, {C3, 3, 1, 1}, {C1, 1, 1, 0}, {C3, 4, 1, 1},
    {C1, 1, 2, 0}, {C1, 2, 0, 0}, {C1, 2, 1, 0}, {C4, 0, 0, 1}, {C1, 2, 2, 0}, {C2, 0, 0, 0}, {C2, 1, 0, 0}, {DN, 0, 0, 1}, {C2, 2, 0, 0}, {C2, 3, 0, 0},
    {C2, 4, 0, 0}};
// Why this order is safe (checked by the eager / lazy host emulation, tests/test_k2_plan_cpu.py):
//   producers >= 4 units before their consumers: C2(r) after the C1 rows 2r..2r+3 it reads; conv3 tile k of the previous batch after
//   that batch's C2 rows (end of the previous iteration); C4' 4 after the last conv3 tile; DN' 4 after C4';
//   an epilogue never overwrites what an older batch's MMAs still read: C2(r0) comes after conv3 tiles 0, 1 (the readers of c2p row
//   0), C2(r1) after tile 3, C2(r2) after tile 4; the conv1 units of batch b + 1 come after every C2 of batch b.
constexpr int kMmaPerIter = 9 * 3 + 5 * 12 + 5 * 9 + 18 + 3;   // 153 MMAs per iteration (checked against mma_count in the host emulation)
// The team that drains the dense unit (one warp's work: lane = clip) also lays out the NEXT batch's input with its other three warps:
// conv1 of this batch is done with c0 by then (all MMAs before the dense unit's have completed), and the slot's next user, four units
// later, is the first conv1 unit of the next iteration -- the one unit that needs the new layout, so its slot wait is the dependency.
// (At position 15, C2(r0), the extra ~2,500 cycles held a slot that C2(r3) was waiting for: profiles/r02g_k2_trace.txt.)
constexpr int kPosTranspose = 17;
constexpr int kPosNextRaw = 0;      // first unit of an iteration, same slot as 17 of the iteration before: the raw buffer has been consumed,
                                    // the issuing thread starts the bulk copy of the batch after the next

static_assert(kSched[kPosTranspose].kind == DN && kUnits - kPosTranspose + kPosNextRaw == kSlots, "transpose slot = first slot of the next iteration");

// ---- incremental sliding-window mode (streaming) ------------------------------------------------------
// A window that starts one input row later is a new window for the pooled layers (the 2x1 pools are anchored at the window
// start), but one that starts FOUR rows later is the old one moved up by exactly one pooled conv2 row and one conv3 row:
//     c2p_h[row 3] = c2p_{h-4}[row 4],  c2p_h[row 2] = c2p_{h-8}[row 4],   c3_h[row 1] = c3_{h-4}[row 2],  c3_h[row 0] = c3_{h-8}[row 2]
// (h = hop number).  So a hop only has to compute the BOTTOM row of every layer -- conv1 rows 16..23 (units C1(2, x)), pooled conv2
// row 4 (C2(4)), conv3 row 2 (tiles 3 and 4), then conv4 and the dense layer -- if the rows it computed 4 and 8 hops ago are
// at hand.  They are kept in HBM per batch of 32 streams in the shared-memory layout, in four phase sets (h mod 4) of three slots
// used in rotation (use k of a set writes slot k % 3 and reads the rows of uses k - 1 and k - 2: a hop never reads a slot it writes).
// The machinery is the same (nnom_q7_umma_incr_kernel, schedule kSchedI below): the epilogues of C2(4) and of the conv3 tiles also store their
// 16-byte results to the cache, and the team of the unit that follows the last reader of a layer's cached rows starts the bulk copies of the next
// batch's rows (contiguous pieces, 34 KB per batch) onto an mbarrier that the issuing thread waits on before the first unit that reads them.
// 60 MMAs per batch instead of 153; bit-identical to the full window (integer convolutions are position-local).
constexpr uint32_t CACHE_C3_ROW = 5 * PX, CACHE_C2_ROW = 7 * PX;                       // one plane of one cached row
constexpr uint32_t CACHE_C3_SLOT = 4 * CACHE_C3_ROW, CACHE_C2_SLOT = 2 * CACHE_C2_ROW; // all planes of a row
constexpr uint32_t CACHE_C2_BASE = 4 * 3 * CACHE_C3_SLOT;                              // conv3 part first: [set 4][slot 3][plane 4][5 px]
constexpr uint32_t CACHE_BATCH_BYTES = CACHE_C2_BASE + 4 * 3 * CACHE_C2_SLOT;          // then conv2: [set 4][slot 3][plane 2][7 px]
constexpr uint32_t kNoCache = 0xffffffffu;
constexpr uint32_t BAR_CACHE = 136;     // mbarrier of conv3's cached-row copies, in the barrier block behind the class histogram
K2_HD uint32_t cache_c3_off(int set, int slot) { return (uint32_t)(set * 3 + slot) * CACHE_C3_SLOT; }
K2_HD uint32_t cache_c2_off(int set, int slot) { return CACHE_C2_BASE + (uint32_t)(set * 3 + slot) * CACHE_C2_SLOT; }

// The incremental mode's own schedule (nnom_q7_umma_incr_kernel): the eight units a hop needs, one pipeline stage per iteration -- in
// iteration i, conv1 works on batch i, conv2 on batch i - 1, conv3 on i - 2, conv4 on i - 3, dense on i - 4 -- and within an iteration
// the deepest stage first.  Every consumer then sits 5..7 units after its producer (previous iteration), and every writer of a
// buffer is issued after the readers of the buffer's previous content (same iteration, earlier position), which is all the slot
// ring needs (see the argument above).  Slots are assigned so that the two long epilogues (dense + input layout, conv2 + its cache traffic) do not
// share one: slot_i below, two uses per slot and iteration, and every producer's slot is re-used (hence waited for) between the producer and its
// consumers.  The next batch's input rows are laid out by the two idle warps of conv3 tile 3's team (its positions 12, 13 are wrap-around pixels),
// whose slot's next user is the first conv1 unit.
constexpr int kUnitsI = 8, kMaxLagI = 4;
constexpr Unit kSchedI[kUnitsI] = {{DN, 0, 0, 4}, {C4, 0, 0, 3}, {C3, 3, 1, 2}, {C3, 4, 1, 2}, {C2, 4, 0, 1}, {C1, 2, 0, 0}, {C1, 2, 1, 0}, {C1, 2, 2, 0}};
constexpr int kPosI_Dense = 0;
constexpr int kPosI_Conv4 = 1;    // the issuing thread waits for conv3's cached rows 0, 1 of its batch here
constexpr int kPosI_Conv3 = 2;    // ... and for conv2's cached rows 2, 3 here; this unit's team starts the copy of conv3's rows for ITS batch and lays out batch i's input
constexpr int kPosI_Conv2 = 4;    // this unit's team starts the copy of conv2's rows for its batch
constexpr int kPosI_Conv1 = 5;    // first conv1 unit (same slot as conv3 tile 3): the issuing thread starts the next raw copy here
// TMEM slot of each position: {0, 1, 2, 3, 1, 2, 3, 0}; positions 0..3 are a slot's first use of an iteration, 4..7 its second
K2_HD constexpr int slot_i(int pos) { return pos < 4 ? pos : (pos == 7 ? 0 : pos - 3); }
// the two positions each slot's team drains, in order: {0, 7}, {1, 4}, {2, 5}, {3, 6}
K2_HD constexpr int team_pos_i(int team, int second) { return second ? (team == 0 ? 7 : team + 3) : team; }
constexpr uint32_t CACHE_C3_LOAD = 2 * CACHE_C3_SLOT, CACHE_C2_LOAD = 2 * CACHE_C2_SLOT;
constexpr uint32_t BAR_CACHE2 = 144;    // second cache mbarrier (conv2 rows); BAR_CACHE takes conv3's
constexpr int kIncrFirstRow = 16;       // conv1 rows 16..23 read input rows 16..27
struct MmaOp { uint32_t a_off, a_lbo, b_off, b_lbo, n, col, acc; };   // SBO = 128 for every operand; col relative to the slot

// Closed forms instead of tables: the kernel's loops must stay small AND must not stream tables through the constant cache
// (sixteen-plus specialised unit bodies were ~170 KB of SASS and every unit's code came from L2 once per iteration; 13 KB of
// per-unit tables in constant memory missed that cache just as steadily: profiles/r02b_k2_trace.txt).
K2_HD constexpr int mma_count(const Unit &u) { return u.kind == C1 ? 3 : u.kind == C2 ? 12 : u.kind == C3 ? 9 * u.b : u.kind == C4 ? 18 : 3; }
K2_HD constexpr uint32_t mma_n(const Unit &u) { return u.kind == C1 ? 96u : (u.kind == C3 || u.kind == C2) ? 64u : u.kind == DN ? 16u : 32u; }

// the m-th MMA of unit u, in issue order
K2_HD constexpr MmaOp mma_op(const Unit &u, int m) {
  if (u.kind == C1) {
    // lanes quarter i <-> conv rows 8t + 2i (tile E, columns 0..47) and 8t + 2i + 1 (tile O, columns 48..95); chunk pair j = m:
    // input rows 2(4t + i + j) and 2(4t + i + j) + 1.  Both tiles read the same A rows, so ONE MMA with N = 96 serves them:
    // B rows 0..47 hold the kernel rows tile E needs from this chunk pair (ky = 2j, 2j + 1), rows 48..95 those of tile O
    // (ky = 2j - 1, 2j), for the three output pixels of x third u.b.
    return MmaOp{OFF_C0 + (uint32_t)(4 * u.a + m) * PX, C0_HALF, OFF_W1 + (uint32_t)m * 2 * W1_CHUNK + (uint32_t)u.b * 96 * 16, W1_CHUNK, 96, 0,
                 (uint32_t)(m > 0)};
  } else if (u.kind == C2) {
    // pooled row r, x half h (pixels x = 4h..4h+3): ONE accumulator tile of N = 64 holds both conv rows under the pooled row --
    // columns 0..31 the odd row y = 2r + 1, columns 32..63 the even row y = 2r (like conv1's E / O tiles: same A rows, so the
    // 2x1 max pool is a max between two registers of one lane).  A walks the four input rows 2r + i, i = 0..3; the B block
    // pair for input row i is [w(ky = i - 1) | w(ky = i)] with zero blocks for ky = -1 and ky = 3, which the block order
    // Z w0 w1 w2 Z provides as 1,024 contiguous bytes starting at block i.  Six MMAs per x half (conv2 used to be five N = 32
    // MMAs per conv row and x half: 20 per unit; an MMA costs the same ~60 cycles for N = 32 and N = 64):
    //   mm = 0..3: input row i = mm, taps kx = 0 (chunk 0) and kx = 1 (chunk 1 = next pixel);
    //   mm = 4, 5: kx = 2 of input rows 2(mm - 4) (chunk 0) and 2(mm - 4) + 1 (chunk 1 = next row)
    const int h = m / 6, mm = m - 6 * h;
    const uint32_t y0 = 2 * u.a, col = (uint32_t)h * 64;
    if (mm < 4) return MmaOp{OFF_C1P + ((y0 + mm) * 9 + 4 * h) * PX, PX, OFF_W2 + (uint32_t)mm * W2_BLOCK, 4 * W2_BLOCK, 64, col, (uint32_t)(mm > 0)};
    const uint32_t i0 = 2 * (uint32_t)(mm - 4);
    return MmaOp{OFF_C1P + ((y0 + i0) * 9 + 4 * h + 2) * PX, 9 * PX, OFF_W2 + (8 + i0) * W2_BLOCK, W2_BLOCK, 64, col, 1};
  } else if (u.kind == C3) {
    // tile tau: output positions q = 4 tau .. 4 tau + 3 on the 7-wide input grid (y = q / 7, x = q % 7; x >= 5 dropped)
    const int tl = m / 9, tap = m - 9 * tl, ky = tap / 3, kx = tap - 3 * ky;
    return MmaOp{OFF_C2P + (uint32_t)(4 * (u.a + tl) + ky * 7 + kx) * PX, PLANE2, OFF_W3 + (uint32_t)tap * 2048, 1024, 64, (uint32_t)tl * 64,
                 (uint32_t)(tap > 0)};
  } else if (u.kind == C4) {
    // one tile: output positions x = 0..3 (3 dropped) on the 5-wide grid; per tap two MMAs (channels 0..31, 32..63)
    const int tap = m >> 1, cp = m & 1, ky = tap / 3, kx = tap - 3 * ky;
    return MmaOp{OFF_C3 + (uint32_t)(ky * 5 + kx) * PX + (uint32_t)cp * 2 * PLANE3, PLANE3, OFF_W4 + (uint32_t)m * 1024, 512, 32, 0, (uint32_t)(m > 0)};
  }
  // dense: rows = clips (lanes 0..31; the other 96 rows read neighbouring pixels and are ignored), K = 3 px x 32 channels
  return MmaOp{OFF_C4 + (uint32_t)m * PX, PLANE4, OFF_WD + (uint32_t)m * 512, 256, 16, 0, (uint32_t)(m > 0)};
}

K2_HD constexpr uint32_t desc_lo(uint32_t off, uint32_t lbo) { return ((off >> 4) & 0x3fffu) | (((lbo >> 4) & 0x3fffu) << 16); }
// cute::UMMA::InstrDescriptor for kind::i8: D = S32 (2 << 4), A = B = signed 8 bit (1 << 7, 1 << 10), K-major, N >> 3 at 17, M >> 4 at 24
K2_HD constexpr uint32_t idesc_i8(uint32_t n) { return (2u << 4) | (1u << 7) | (1u << 10) | ((n >> 3) << 17) | ((128u >> 4) << 24); }

// ---- quantisation epilogue -------------------------------------------------------------------------------------------
// clamp((m + seed) >> rs, 0, 127) = (max(min(m + seed, hi), 0)) >> rs with hi = (128 << rs) - 1: SSAT8 then ReLU
// (nnom_activation.c:118-126) of a conv output whose seed already holds (bias << lshift) + round (nnom_local.c:425-484)
K2_HD int q7_relu(int m, int seed, int hi, int rs) {
#if defined(__CUDA_ARCH__)
  return __viaddmin_s32_relu(m, seed, hi) >> rs;
#else
  int v = m + seed;
  v = v < hi ? v : hi;
  v = v > 0 ? v : 0;
  return v >> rs;
#endif
}
K2_HD uint32_t pack4(int a, int b, int c, int d) { return (uint32_t)a | ((uint32_t)b << 8) | ((uint32_t)c << 16) | ((uint32_t)d << 24); }
K2_HD int imax(int a, int b) { return a > b ? a : b; }

struct Vec16 { uint32_t x, y, z, w; };
struct Int4 { int x, y, z, w; };

// 16 consecutive int32 parameters starting at index i (a multiple of 4) as four 128-bit broadcast loads
template <class Ctx> K2_HD void load16(Ctx &c, int i, int (&s)[16]) {
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const Int4 v = c.par4(i + 4 * k);
    s[4 * k] = v.x; s[4 * k + 1] = v.y; s[4 * k + 2] = v.z; s[4 * k + 3] = v.w;
  }
}

// 16 accumulators (optionally the max with 16 more: the 2x1 pool) -> 16 int8 activations
template <bool POOL> K2_HD Vec16 quant16(const uint32_t (&e)[16], const uint32_t (&o)[16], const int *seed, int hi, int rs) {
  int v[16];
#pragma unroll
  for (int i = 0; i < 16; i++) v[i] = q7_relu(POOL ? imax((int)e[i], (int)o[i]) : (int)e[i], seed[i], hi, rs);
  return Vec16{pack4(v[0], v[1], v[2], v[3]), pack4(v[4], v[5], v[6], v[7]), pack4(v[8], v[9], v[10], v[11]),
               pack4(v[12], v[13], v[14], v[15])};
}

K2_HD int ssat8(int v) { return v < -128 ? -128 : (v > 127 ? 127 : v); }
K2_HD int usat(int v, int bits) { return v < 0 ? 0 : (v > (1 << bits) - 1 ? (1 << bits) - 1 : v); }

// arm_softmax_q7 (reference branch arm_softmax_q7.c:217-254, SIMD branch :83-215) and local_softmax_q7 (nnom_local.c:948-1003)
// mode: EDISON_SOFTMAX_* (0 CMSIS reference, 1 CMSIS SIMD with 8-bit wrapping subtractions, 2 NNoM local)
// (every loop is unrolled: in and out are register arrays in the kernel -- as rolled loops they lived in local memory, on the one warp
// whose latency the dense unit's slot waits for)
K2_HD void softmax_q7(const int8_t *in, int8_t *out, int mode) {
  constexpr int n = 10;
  if (mode == 2) {
    int base = -257;
#pragma unroll
    for (int i = 0; i < n; i++) base = imax(base, (int)in[i]);
    base -= 8;
    int sum = 0;
#pragma unroll
    for (int i = 0; i < n; i++)
      if (in[i] > base) sum += 1 << usat(in[i] - base, 5);
    const int ob = 0x100000 / sum;
#pragma unroll
    for (int i = 0; i < n; i++) out[i] = in[i] > base ? (int8_t)ssat8(ob >> usat(13 + base - in[i], 5)) : (int8_t)0;
    return;
  }
  int base = -128;
#pragma unroll
  for (int i = 0; i < n; i++) base = imax(base, (int)in[i]);
  base -= 8;
  const int nv = mode == 1 ? (n & ~3) : 0;   // handled 4 at a time with 8-bit wrap
  int sum = 0;
#pragma unroll
  for (int i = 0; i < n; i++) {
    int d = in[i] - base;
    if (i < nv) d = (int)(int8_t)(d & 0xff);   // __SSUB8 wraps (:127)
    sum += 1 << usat(d, 3);
  }
  const int ob = (1 << 20) / sum;
#pragma unroll
  for (int i = 0; i < n; i++) {
    int d;
    if (i < nv) {
      const int t = (int)(int8_t)((13 - in[i]) & 0xff);               // __SSUB8(pad, in) :177
      d = (int)(int8_t)((t + (int)(int8_t)(base & 0xff)) & 0xff);     // __SADD8(in, baseV) :178
    } else {
      d = 13 + base - in[i];
    }
    out[i] = (int8_t)ssat8(ob >> usat(d, 5));
  }
}

// What an epilogue needs from its executor (CUDA kernel or host emulation):
//   c.ld16(col, v)   : this thread's lane of TMEM columns col..col+15 of the unit's slot;  c.ld_wait() before using v
//   c.st16(off, Vec16) : 16-byte shared-memory store;  c.st16_cache(off, Vec16): 16-byte store into this batch's row cache (incremental mode:
//   c.cache_w_c2 / c.cache_w_c3 = offsets of the slots being written, kNoCache otherwise);  c.par(i) / c.par4(i): int32 parameter(s) i (same for all lanes)
//   c.lane (= clip), c.q (lane quarter = pixel within an M = 128 tile);  dense only: c.clip0 (first clip of the batch),
//   c.nclips, c.logits, c.softmax (global, may be null), c.softmax_mode, c.counting + c.count(cls) (class histogram hook),
//   c.filt (null or the streaming filter state) with c.prefetch(clip) (start loading the state) and c.post(clip, softmax)
// One warp per lane quarter drains a whole unit: lean loops (a job = 16 accumulator columns -> one 16-byte store is ~75
// instructions; evaluating a generic job descriptor per job tripled that and the epilogue is latency bound).
struct JobParams { uint32_t col_e, col_o, seed, out, gout; };   // gout: byte offset in this batch's row cache, or kNoCache

// n jobs back to back, software pipelined: the tcgen05.ld of job j + 1 is in flight while job j is quantised and stored (the
// epilogue is bound by latency -- ~250 cycles of TMEM load against ~100 of arithmetic per job -- not by issue slots)
template <bool POOL, class Ctx, class P> K2_HD void job_stream(Ctx &c, int n, int rs, int hi, P params) {
  if (n <= 0) return;
  uint32_t e[16], o[16];
  JobParams p = params(0);
  c.ld16(p.col_e, e);
  if (POOL) c.ld16(p.col_o, o);
#pragma unroll 1
  for (int j = 0; j < n; j++) {
    int seed[16];
    load16(c, (int)p.seed, seed);   // in flight with the tcgen05.ld
    c.ld_wait();
    int m[16];
#pragma unroll
    for (int i = 0; i < 16; i++) m[i] = POOL ? imax((int)e[i], (int)o[i]) : (int)e[i];
    const JobParams cur = p;
    if (j + 1 < n) {
      p = params(j + 1);
      c.ld16(p.col_e, e);
      if (POOL) c.ld16(p.col_o, o);
    }
    int v[16];
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = q7_relu(m[i], seed[i], hi, rs);
    const Vec16 r{pack4(v[0], v[1], v[2], v[3]), pack4(v[4], v[5], v[6], v[7]), pack4(v[8], v[9], v[10], v[11]), pack4(v[12], v[13], v[14], v[15])};
    c.st16(cur.out, r);
    if (cur.gout != kNoCache) c.st16_cache(cur.gout, r);
  }
}

template <class Ctx> K2_HD void epilogue_conv(Ctx &c, const Unit &u) {
  const uint32_t lane16 = (uint32_t)c.lane * 16;
  if (u.kind == C1) {
    // lane quarter q is pooled row 4t + q; the unit holds the three pixels of x third u.b: tile E at columns 16 px, tile O at 48 + 16 px
    const uint32_t out = OFF_C1P + (uint32_t)((4 * u.a + c.q) * 9 + 3 * u.b) * PX + lane16;
    job_stream<true>(c, 3, c.par(PAR_RS + 0), c.par(PAR_HI + 0), [&](int j) {
      return JobParams{(uint32_t)j * 16, 48 + (uint32_t)j * 16, (uint32_t)PAR_SEED1, out + (uint32_t)j * PX, kNoCache};
    });
  } else if (u.kind == C2) {
    // pooled row u.a; job j: x half h = j >> 1, channel half j & 1; x = 4h + q (x = 7, i.e. q = 3 in the second half, is the
    // wrap-around pixel); x half h at columns 64 h: odd conv row at +0, even conv row at +32
    const uint32_t out = OFF_C2P + (uint32_t)(u.a * 7 + c.q) * PX + lane16;
    // incremental mode: pooled row 4 also goes to the row cache, [plane][x][clip][16]
    const uint32_t gbase = (c.cache_w_c2 != kNoCache && u.a == 4) ? c.cache_w_c2 + (uint32_t)c.q * PX + lane16 : kNoCache;
    job_stream<true>(c, c.q < 3 ? 4 : 2, c.par(PAR_RS + 1), c.par(PAR_HI + 1), [&](int j) {
      const uint32_t h = (uint32_t)j >> 1, half = (uint32_t)j & 1;
      return JobParams{h * 64 + 32 + half * 16, h * 64 + half * 16, (uint32_t)PAR_SEED2 + half * 16, out + half * PLANE2 + h * 4 * PX,
                       gbase == kNoCache ? kNoCache : gbase + half * CACHE_C2_ROW + h * 4 * PX};
    });
  } else if (u.kind == C3) {
    // one tile: position idx = 4 u.a + q on the 7-wide grid (x >= 5 and row 3 are wrap-around positions); four 16-channel planes
    const int idx = 4 * u.a + c.q, y = idx / 7, x = idx - 7 * y;
    const uint32_t out = OFF_C3 + (uint32_t)(y * 5 + x) * PX + lane16;
    // incremental mode: row 2 also goes to the row cache, [plane][x][clip][16]
    const uint32_t gbase = (c.cache_w_c3 != kNoCache && y == 2) ? c.cache_w_c3 + (uint32_t)x * PX + lane16 : kNoCache;
    job_stream<false>(c, (x < 5 && y < 3) ? 4 : 0, c.par(PAR_RS + 2), c.par(PAR_HI + 2), [&](int j) {
      return JobParams{(uint32_t)j * 16, 0, (uint32_t)PAR_SEED3 + (uint32_t)j * 16, out + (uint32_t)j * PLANE3,
                       gbase == kNoCache ? kNoCache : gbase + (uint32_t)j * CACHE_C3_ROW};
    });
  } else if (u.kind == C4) {
    // x = q (3 is the wrap-around pixel); two 16-channel planes
    const uint32_t out = OFF_C4 + (uint32_t)c.q * PX + lane16;
    job_stream<false>(c, c.q < 3 ? 2 : 0, c.par(PAR_RS + 3), c.par(PAR_HI + 3), [&](int j) {
      return JobParams{(uint32_t)j * 16, 0, (uint32_t)PAR_SEED4 + (uint32_t)j * 16, out + (uint32_t)j * PLANE4, kNoCache};
    });
  }
}

// Dense 96 -> 10 (nnom_dense.c:92-119): lane = clip (lane quarter 0 only); logits = SSAT8((acc + seed) >> rs); then softmax
template <class Ctx> K2_HD void run_dense(Ctx &c) {
  const int rs = c.par(PAR_RS + 4);
  uint32_t e[16];
  c.ld16(0, e);
  if (c.filt && c.lane < c.nclips) c.prefetch(c.clip0 + c.lane);   // streaming: the filter state travels while the accumulators do
  c.ld_wait();
  if (c.lane < c.nclips) {
    int8_t lg[10];
    int seed[16];
    load16(c, PAR_SEEDD, seed);
#pragma unroll
    for (int i = 0; i < 10; i++) lg[i] = (int8_t)ssat8(((int)e[i] + seed[i]) >> rs);
    const long long clip = c.clip0 + c.lane;
    if (c.logits)
#pragma unroll
      for (int i = 0; i < 10; i++) c.logits[clip * 10 + i] = lg[i];
    if (c.softmax || c.filt) {
      int8_t sm[10];
      softmax_q7(lg, sm, c.softmax_mode);
      if (c.softmax)
#pragma unroll
        for (int i = 0; i < 10; i++) c.softmax[clip * 10 + i] = sm[i];
      if (c.filt) c.post(clip, sm);   // streaming: moving average, argmax, threshold (app.c:336-350)
    }
    if (c.counting) {   // class histogram (edison_kws_counts_dev): first maximum of the logits, the class the firmware's arm_max would report on them
      int best = 0, bestv = lg[0];   // (the value is tracked beside the index: lg[best] would be a run-time index into a register array)
#pragma unroll
      for (int i = 1; i < 10; i++)
        if (lg[i] > bestv) { bestv = lg[i]; best = i; }
      c.count(best);
    }
  }
}

// input layout step: the 13 bytes of input row `row` of clip `clip` (raw = [clip][403]) -> c0[row & 1][row >> 1][clip][16].
// w[0..3] are the four aligned 32-bit words that hold raw bytes a..a+12, a = clip * 403 + row * 13; bytes 13..15 of the chunk
// only ever meet zero weights.
K2_HD Vec16 input_chunk(const uint32_t (&w)[4], uint32_t a) {
  const uint32_t sh = (a & 3) * 8;
  Vec16 v;
  if (sh == 0) return Vec16{w[0], w[1], w[2], w[3]};
  v.x = (w[0] >> sh) | (w[1] << (32 - sh));
  v.y = (w[1] >> sh) | (w[2] << (32 - sh));
  v.z = (w[2] >> sh) | (w[3] << (32 - sh));
  v.w = w[3] >> sh;
  return v;
}
// byte offset of input row `row` of clip `clip` in raw [clip][403]; windows of the streaming path are rings of rows: logical row r
// lives at physical row (r + row0) % 31
K2_HD uint32_t raw_offset(int clip, int row, int row0) {
  const int phys = row + row0 >= 31 ? row + row0 - 31 : row + row0;
  return (uint32_t)clip * kNetIn + (uint32_t)phys * 13;
}
K2_HD uint32_t c0_offset(int row, int clip) { return OFF_C0 + (uint32_t)(row & 1) * C0_HALF + (uint32_t)(row >> 1) * PX + (uint32_t)clip * 16; }
constexpr int kInRowsUsed = 28;   // conv rows 24..26 only feed the pooled row that conv2's pooling drops (and row 26 nothing)

// ---- weight image (host) ---------------------------------------------------------------------------------------------
// w[l], b[l]: the five layers' int8 weights in the reference's O,H,W,I order and biases; rs / ls: output right shifts and bias
// left shifts (weights.h:72-105).  dense weights already de-interleaved to [10][96].
inline void build_image(uint8_t *img, const int8_t *const w[5], const int8_t *const b[5], const int rs[5], const int ls[5]) {
  memset(img, 0, IMAGE_BYTES);
  // conv1: chunk (pair j, chunk h) = 288 rows x 16 bytes; row = xs * 96 + p * 48 + px * 16 + c for output pixel x = 3 xs + px of
  // tile p (0 = even conv rows, 1 = odd); byte kk = input x; the row holds kernel row ky = 2j + h (E) or 2j - 1 + h (O):
  //   w1[c][ky][kk - x] for 0 <= kk - x < 5, zero elsewhere and for ky outside 0..4
  for (int j = 0; j < 3; j++)
    for (int h = 0; h < 2; h++)
      for (int p = 0; p < 2; p++) {
        const int ky = p ? 2 * j - 1 + h : 2 * j + h;
        if (ky < 0 || ky > 4) continue;
        uint8_t *chunk = img + OFF_W1 + (j * 2 + h) * W1_CHUNK;
        for (int x = 0; x < 9; x++)
          for (int c = 0; c < 16; c++)
            for (int kx = 0; kx < 5; kx++)
              chunk[((x / 3) * 96 + p * 48 + (x % 3) * 16 + c) * 16 + x + kx] = (uint8_t)w[0][(c * 5 + ky) * 5 + kx];
      }
  // conv2: block 4 kx + 1 + ky = [n][16] weights of tap (ky, kx); blocks 0, 4, 8, 12 stay zero (mma_op, C2)
  for (int kx = 0; kx < 3; kx++)
    for (int ky = 0; ky < 3; ky++)
      for (int n = 0; n < 32; n++)
        for (int kk = 0; kk < 16; kk++)
          img[OFF_W2 + (4 * kx + 1 + ky) * W2_BLOCK + n * 16 + kk] = (uint8_t)w[1][n * 144 + (ky * 3 + kx) * 16 + kk];
  for (int tap = 0; tap < 9; tap++)
    for (int h = 0; h < 2; h++)
      for (int n = 0; n < 64; n++)
        for (int kk = 0; kk < 16; kk++) img[OFF_W3 + tap * 2048 + h * 1024 + n * 16 + kk] = (uint8_t)w[2][n * 288 + tap * 32 + 16 * h + kk];
  for (int tap = 0; tap < 9; tap++)
    for (int cp = 0; cp < 2; cp++)
      for (int h = 0; h < 2; h++)
        for (int n = 0; n < 32; n++)
          for (int kk = 0; kk < 16; kk++)
            img[OFF_W4 + (tap * 2 + cp) * 1024 + h * 512 + n * 16 + kk] = (uint8_t)w[3][n * 576 + tap * 64 + 32 * cp + 16 * h + kk];
  for (int m = 0; m < 3; m++)
    for (int h = 0; h < 2; h++)
      for (int n = 0; n < 10; n++)
        for (int kk = 0; kk < 16; kk++) img[OFF_WD + m * 512 + h * 256 + n * 16 + kk] = (uint8_t)w[4][n * 96 + m * 32 + 16 * h + kk];
  int32_t par[PAR_BYTES / 4] = {0};
  const int seed_at[5] = {PAR_SEED1, PAR_SEED2, PAR_SEED3, PAR_SEED4, PAR_SEEDD}, cout[5] = {16, 32, 64, 32, 10};
  for (int l = 0; l < 5; l++) {
    for (int o = 0; o < cout[l]; o++) par[seed_at[l] + o] = (int32_t)((uint32_t)(int32_t)b[l][o] << ls[l]) + (1 << (rs[l] - 1));
    par[PAR_RS + l] = rs[l];
    par[PAR_HI + l] = (128 << rs[l]) - 1;
  }
  memcpy(img + OFF_PAR, par, PAR_BYTES);
}

}  // namespace k2
}  // namespace edison
