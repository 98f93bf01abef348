// nnom_generic.cu -- generic executor for NNoM int8 models.
//
// K2 (nnom_q7_kernel.cu) is specialised for the reference's kws model.  Any other model that nnom_utils.py generates from
// the vocabulary below is converted to a blob -- tools/convert_nnom_weights.py (version 1: the conv/dense chain of the
// reference model family) or tools/convert_nnom_graph.py (version 2: a graph) -- and runs here node by node with activations
// in HBM: one thread per output element, dp4a over the channel axis when it is a multiple of 4.
//   Input | Conv2D any kernel / stride, PADDING_VALID or PADDING_SAME (+ ReLU tail) | MaxPool any kernel / stride / padding |
//   GlobalMaxPool | Concat on the channel axis | Add with output shift (+ ReLU tail) | Dense (+ ReLU tail) | Softmax
// Same arithmetic as the reference's NNoM runtime, local backend (firmware/src/ai/nnom/src):
//   conv nnom_conv2d.c:77-82, :101-112 + nnom_local.c:425-483; max pool nnom_maxpool.c:66-104 + nnom_local.c:117-159;
//   global max pool nnom_global_pool.c:93-96; concat nnom_concat.c:186-213; add nnom_local.c:1109-1126;
//   dense nnom_local.c:804-921; softmax variants as in nnom_q7_kernel.cu
// so results are bit-exact; it is a correctness-first path, not a tuned kernel.  Pinned against three upstream NNoM example
// models inside the reference tree (keyword_spotting, mnist-simple, uci-inception) run through the reference's own runtime
// built for the host (tests/golden/nnom_example.npz, nnom_graphs.npz).  Not supported, because this NNoM version does not
// compute them reproducibly: Concat of unequal inputs (nnom_concat.c:104, :208 size every input like the first) and
// UpSample (nnom_local.c:383 never writes the odd rows).
#include <stdio.h>
#include <string.h>

#include <vector>

#include "edison_internal.h"
#include "../../include/edison_b200.h"

namespace edison {

enum { G_INPUT, G_CONV, G_DENSE, G_MAXPOOL, G_CONCAT, G_SOFTMAX, G_GMAXPOOL, G_ADD };
constexpr int kMaxNodes = 64;

struct GenNode {
  int op, n_in, in[8];
  int cout, kh, kw, sh, sw, ph, pw, same, rshift, lshift, relu;
  int H, W, C;                        // output shape
  int aH, aW, aC;                     // shape of the first input
  long long w_off, b_off;             // into the device weight buffer
  long long act_off, act_size;        // per input: offset and size of this node's output in the arena
};

__device__ __forceinline__ int g_ssat8(int v) { return min(max(v, -128), 127); }
__device__ __forceinline__ int g_usat(int v, int bits) { return min(max(v, 0), (1 << bits) - 1); }

// conv + optional ReLU.  in: [n][aH][aW][aC], out: [n][H][W][C]; taps outside the image are skipped
__global__ void gen_conv_kernel(const int8_t *__restrict__ in, int8_t *__restrict__ out, const int8_t *__restrict__ w,
                                const int8_t *__restrict__ b, GenNode L, long long n) {
  const int O = L.C, C = L.aC;
  const long long total = n * L.H * L.W * O;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int o = (int)(idx % O);
    long long r = idx / O;
    const int x = (int)(r % L.W); r /= L.W;
    const int y = (int)(r % L.H);
    const long long clip = r / L.H;
    const int8_t *src = in + clip * ((long long)L.aH * L.aW * C);
    const int8_t *wo = w + (long long)o * L.kh * L.kw * C;
    int acc = (int)((unsigned)(int)b[o] << L.lshift) + (1 << (L.rshift - 1));
    const int c0 = x * L.sw - L.pw;
    const int kx0 = c0 < 0 ? -c0 : 0, kx1 = min(L.kw, L.aW - c0);   // the valid taps of a row are contiguous in both operands
    for (int ky = 0; ky < L.kh; ky++) {
      const int row = y * L.sh - L.ph + ky;
      if (row < 0 || row >= L.aH || kx1 <= kx0) continue;
      const int8_t *a = src + ((long long)row * L.aW + c0 + kx0) * C;
      const int8_t *wr = wo + (ky * L.kw + kx0) * C;
      const int len = (kx1 - kx0) * C;
      int i = 0;
      if ((C & 3) == 0) {
        for (; i < len; i += 4)
          acc = __dp4a(*reinterpret_cast<const int *>(a + i), *reinterpret_cast<const int *>(wr + i), acc);
      } else {
        for (; i < len; i++) acc += (int)a[i] * (int)wr[i];
      }
    }
    int v = g_ssat8(acc >> L.rshift);
    if (L.relu && v < 0) v = 0;
    out[idx] = (int8_t)v;
  }
}

// ---- tensor-core path for conv / dense nodes whose input channel count is a multiple of 4 ------------------------------------
// Implicit GEMM on mma.sync.m16n8k32.s8 (the per-node activations live in HBM here, so operands are register fragments, not
// shared-memory descriptors: K2 is the tcgen05 path of the tuned model): M = output pixels of all inputs, N = output channels,
// K = (ky, kx, c) in the reference's O,H,W,I weight order.  With C % 4 == 0 every A register (4 consecutive k) is 4 channels of
// one tap = one aligned 32-bit load from the [n][H][W][C] activation buffer; taps outside the image contribute zero, which is
// what the reference's tap skipping computes (nnom_local.c:425-483).  A warp owns 32 output pixels x up to 8 NT output
// channels; the (tap -> byte offset) table of the node is built once per CTA in shared memory.  int32 accumulation is exact and
// order independent, the epilogue is the reference's (bias << lshift) + round, >> rshift, SSAT8, ReLU: bit-exact.
__device__ __forceinline__ void gen_mma_s8(int (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

constexpr int kGenMmaWarps = 8;

// BYTES (input channels not a multiple of 4, e.g. the single-channel first layer of every model here): the table has one entry per
// k and a register is assembled from four byte loads, each with its own tap test; weights likewise.  Still ~25x fewer loads per
// output than one thread per output element.
template <int NT, bool BYTES>
__global__ void __launch_bounds__(kGenMmaWarps * 32) gen_conv_mma_kernel(const int8_t *__restrict__ in, int8_t *__restrict__ out,
                                                                         const int8_t *__restrict__ w, const int8_t *__restrict__ b,
                                                                         GenNode L, long long n_rows, int K, int bounds) {
  extern __shared__ int2 s_tab[];   // per group of 4 k: x = byte offset from the pixel's window origin, y = (ky << 16) | kx, -1 past K
  const int C = L.aC, O = L.C, HW = L.H * L.W;
  const int step = BYTES ? 1 : 4;                                   // k per table entry
  const int nq = BYTES ? K : K >> 2, nq_pad = BYTES ? (K + 31) & ~31 : (nq + 7) & ~7;
  for (int q = threadIdx.x; q < nq_pad; q += blockDim.x) {
    int2 e = make_int2(0, -1);
    if (q < nq) {
      const int k = step * q, tap = k / C, c = k - tap * C, ky = tap / L.kw, kx = tap - ky * L.kw;
      e = make_int2((ky * L.aW + kx) * C + c, (ky << 16) | kx);
    }
    s_tab[q] = e;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  const long long n_tiles = (n_rows + 31) >> 5;
  for (long long tile = (long long)blockIdx.x * kGenMmaWarps + warp; tile < n_tiles; tile += (long long)gridDim.x * kGenMmaWarps) {
    // the four output pixels this thread feeds: rows 16 mt + 8 h + g of the tile
    const int8_t *base[2][2];
    int y0[2][2], x0[2][2];
    long long row[2][2];
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const long long r = tile * 32 + mt * 16 + h * 8 + g;
        row[mt][h] = r;
        const long long rr = r < n_rows ? r : n_rows - 1;   // rows past the end recompute the last one and are not stored
        const long long clip = rr / HW;
        const int rem = (int)(rr - clip * HW), y = rem / L.W, x = rem - y * L.W;
        y0[mt][h] = y * L.sh - L.ph;
        x0[mt][h] = x * L.sw - L.pw;
        base[mt][h] = in + clip * ((long long)L.aH * L.aW * C) + ((long long)y0[mt][h] * L.aW + x0[mt][h]) * C;
      }
    auto fetch = [&](int mt, int h, int2 e) -> uint32_t {
      if (e.y < 0) return 0u;
      if (bounds) {
        const int yy = y0[mt][h] + (e.y >> 16), xx = x0[mt][h] + (e.y & 0xffff);
        if (yy < 0 || yy >= L.aH || xx < 0 || xx >= L.aW) return 0u;
      }
      return __ldg(reinterpret_cast<const unsigned int *>(base[mt][h] + e.x));
    };
    auto fetch_bytes = [&](int mt, int h, int k) -> uint32_t {      // k..k+3, one table entry and one byte load each
      uint32_t v = 0;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int2 e = s_tab[k + j];
        if (e.y < 0) continue;
        if (bounds) {
          const int yy = y0[mt][h] + (e.y >> 16), xx = x0[mt][h] + (e.y & 0xffff);
          if (yy < 0 || yy >= L.aH || xx < 0 || xx >= L.aW) continue;
        }
        v |= (uint32_t)(uint8_t)__ldg(base[mt][h] + e.x) << (8 * j);
      }
      return v;
    };
    auto weight_bytes = [&](const int8_t *wr, int k) -> uint32_t {
      uint32_t v = 0;
#pragma unroll
      for (int j = 0; j < 4; j++)
        if (k + j < K) v |= (uint32_t)(uint8_t)__ldg(wr + k + j) << (8 * j);
      return v;
    };
    for (int n0 = 0; n0 < O; n0 += 8 * NT) {
      int acc[2][NT][4];
#pragma unroll
      for (int nt = 0; nt < NT; nt++) {
        const int c0 = n0 + 8 * nt + 2 * t;
        const int s0 = c0 < O ? (int)((unsigned)(int)b[c0] << L.lshift) + (1 << (L.rshift - 1)) : 0;
        const int s1 = c0 + 1 < O ? (int)((unsigned)(int)b[c0 + 1] << L.lshift) + (1 << (L.rshift - 1)) : 0;
#pragma unroll
        for (int mt = 0; mt < 2; mt++) { acc[mt][nt][0] = s0; acc[mt][nt][1] = s1; acc[mt][nt][2] = s0; acc[mt][nt][3] = s1; }
      }
      for (int k0 = 0; k0 < K; k0 += 32) {
        const int klo = k0 + 4 * t, khi = klo + 16;
        uint32_t a[2][4];
        if (BYTES) {
#pragma unroll
          for (int mt = 0; mt < 2; mt++) {
            a[mt][0] = fetch_bytes(mt, 0, klo); a[mt][1] = fetch_bytes(mt, 1, klo);
            a[mt][2] = fetch_bytes(mt, 0, khi); a[mt][3] = fetch_bytes(mt, 1, khi);
          }
        } else {
          const int2 e0 = s_tab[(k0 >> 2) + t], e1 = s_tab[(k0 >> 2) + 4 + t];
#pragma unroll
          for (int mt = 0; mt < 2; mt++) {
            a[mt][0] = fetch(mt, 0, e0); a[mt][1] = fetch(mt, 1, e0);
            a[mt][2] = fetch(mt, 0, e1); a[mt][3] = fetch(mt, 1, e1);
          }
        }
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
          const int col = n0 + 8 * nt + g;
          const int8_t *wr = w + (long long)col * K;
          uint32_t b0, b1;
          if (BYTES) {
            b0 = col < O ? weight_bytes(wr, klo) : 0u;
            b1 = col < O ? weight_bytes(wr, khi) : 0u;
          } else {
            b0 = (col < O && klo < K) ? __ldg(reinterpret_cast<const unsigned int *>(wr + klo)) : 0u;
            b1 = (col < O && khi < K) ? __ldg(reinterpret_cast<const unsigned int *>(wr + khi)) : 0u;
          }
          gen_mma_s8(acc[0][nt], a[0], b0, b1);
          gen_mma_s8(acc[1][nt], a[1], b0, b1);
        }
      }
#pragma unroll
      for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int h = 0; h < 2; h++) {
          if (row[mt][h] >= n_rows) continue;
          int8_t *dst = out + row[mt][h] * O;
#pragma unroll
          for (int nt = 0; nt < NT; nt++)
#pragma unroll
            for (int j = 0; j < 2; j++) {
              const int col = n0 + 8 * nt + 2 * t + j;
              if (col >= O) continue;
              int v = g_ssat8(acc[mt][nt][2 * h + j] >> L.rshift);
              if (L.relu && v < 0) v = 0;
              dst[col] = (int8_t)v;
            }
        }
    }
  }
}

// dense: in [n][cin] (the flattened H,W,C output of the previous node), w [cout][cin]
__global__ void gen_dense_kernel(const int8_t *__restrict__ in, int8_t *__restrict__ out, const int8_t *__restrict__ w,
                                 const int8_t *__restrict__ b, GenNode L, long long n) {
  const int cin = L.aH * L.aW * L.aC;
  const long long total = n * L.C;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int o = (int)(idx % L.C);
    const long long clip = idx / L.C;
    const int8_t *x = in + clip * cin, *wr = w + (long long)o * cin;
    int acc = (int)((unsigned)(int)b[o] << L.lshift) + (1 << (L.rshift - 1));
    int i = 0;
    if ((cin & 3) == 0)
      for (; i < cin; i += 4) acc = __dp4a(*reinterpret_cast<const int *>(x + i), *reinterpret_cast<const int *>(wr + i), acc);
    else
      for (; i < cin; i++) acc += (int)x[i] * (int)wr[i];
    int v = g_ssat8(acc >> L.rshift);
    if (L.relu && v < 0) v = 0;
    out[idx] = (int8_t)v;
  }
}

// max pool (also the global one: kernel = the whole map).  Starts from -129 and stores the int as q7, as the reference does
__global__ void gen_maxpool_kernel(const int8_t *__restrict__ in, int8_t *__restrict__ out, GenNode L, long long n) {
  const int C = L.C;
  const long long total = n * L.H * L.W * C;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(idx % C);
    long long r = idx / C;
    const int x = (int)(r % L.W); r /= L.W;
    const int y = (int)(r % L.H);
    const long long clip = r / L.H;
    const int8_t *src = in + clip * ((long long)L.aH * L.aW * C) + c;
    int mx = -129;
    for (int ky = y * L.sh - L.ph; ky < y * L.sh - L.ph + L.kh; ky++)
      for (int kx = x * L.sw - L.pw; kx < x * L.sw - L.pw + L.kw; kx++)
        if (ky >= 0 && kx >= 0 && ky < L.aH && kx < L.aW) mx = max(mx, (int)src[((long long)ky * L.aW + kx) * C]);
    out[idx] = (int8_t)mx;
  }
}

struct ConcatSrc { const int8_t *p[8]; };

// channel-axis concat of n_in equally shaped inputs
__global__ void gen_concat_kernel(ConcatSrc src, int8_t *__restrict__ out, GenNode L, long long n) {
  const int Ca = L.aC, C = L.C;
  const long long total = n * L.H * L.W * C;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(idx % C);
    const long long px = idx / C;
    const int k = c / Ca;
    out[idx] = src.p[k][px * Ca + (c - k * Ca)];
  }
}

// add with rounding output shift + optional ReLU
__global__ void gen_add_kernel(const int8_t *__restrict__ a, const int8_t *__restrict__ b, int8_t *__restrict__ out, GenNode L,
                               long long n) {
  const long long total = n * L.H * L.W * L.C;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    int v = g_ssat8(((int)a[idx] + (int)b[idx] + (1 << (L.rshift - 1))) >> L.rshift);
    if (L.relu && v < 0) v = 0;
    out[idx] = (int8_t)v;
  }
}

// arm_softmax_q7 (reference :217-254, SIMD :83-215) and local_softmax_q7 (nnom_local.c:948-1003), any vector length
__global__ void gen_softmax_kernel(const int8_t *__restrict__ in, int8_t *__restrict__ out, int nc, long long n, int mode) {
  const long long clip = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (clip >= n) return;
  const int8_t *x = in + clip * nc;
  int8_t *y = out + clip * nc;
  if (mode == EDISON_SOFTMAX_NNOM_LOCAL) {
    int base = -257;
    for (int i = 0; i < nc; i++) base = max(base, (int)x[i]);
    base -= 8;
    int sum = 0;
    for (int i = 0; i < nc; i++)
      if (x[i] > base) sum += 1 << g_usat(x[i] - base, 5);
    const int ob = 0x100000 / sum;
    for (int i = 0; i < nc; i++) y[i] = x[i] > base ? (int8_t)g_ssat8(ob >> g_usat(13 + base - x[i], 5)) : (int8_t)0;
    return;
  }
  int base = -128;
  for (int i = 0; i < nc; i++) base = max(base, (int)x[i]);
  base -= 8;
  const int nv = mode == EDISON_SOFTMAX_CMSIS_SIMD ? (nc & ~3) : 0;
  int sum = 0;
  for (int i = 0; i < nc; i++) {
    int d = x[i] - base;
    if (i < nv) d = (int)(int8_t)(d & 0xff);
    sum += 1 << g_usat(d, 3);
  }
  const int ob = (1 << 20) / sum;
  for (int i = 0; i < nc; i++) {
    int d;
    if (i < nv) {
      const int t = (int)(int8_t)((13 - x[i]) & 0xff);
      d = (int)(int8_t)((t + (int)(int8_t)(base & 0xff)) & 0xff);
    } else {
      d = 13 + base - x[i];
    }
    y[i] = (int8_t)g_ssat8(ob >> g_usat(d, 5));
  }
}

}  // namespace edison

struct edison_model {
  int n_nodes, out_node;
  edison::GenNode N[edison::kMaxNodes];
  int8_t *d_w;               // all weights and biases, 4-byte aligned per array
  int in_h, in_w, in_c, in_size, n_out;
  long long arena_per_input; // bytes per input of all intermediate activations
  int8_t *d_arena, *d_in, *d_lg, *d_sm;
  size_t cap;                // inputs the scratch buffers hold
  int naive;                 // debug: 1 = one thread per output element for every node (edison_model_set_naive)
};

namespace edison {

static int model_reserve(edison_model *m, size_t n) {
  if (n <= m->cap) return EDISON_OK;
  int8_t **bufs[] = {&m->d_arena, &m->d_in, &m->d_lg, &m->d_sm};
  const size_t per[] = {(size_t)m->arena_per_input, (size_t)m->in_size, (size_t)m->n_out, (size_t)m->n_out};
  for (int i = 0; i < 4; i++) {
    if (*bufs[i]) cudaFree(*bufs[i]);
    *bufs[i] = nullptr;
    cudaError_t e = cudaMalloc(bufs[i], n * per[i] + 16);
    if (e != cudaSuccess) { m->cap = 0; return ctx_fail(EDISON_E_NOMEM, "edison_model scratch for %zu inputs: %s", n, cudaGetErrorString(e)); }
  }
  m->cap = n;
  return EDISON_OK;
}

static int model_run(edison_model *m, const CtxView &c, const int8_t *d_in, size_t n, int8_t *d_logits, int8_t *d_softmax,
                     cudaStream_t st) {
  const int threads = 256;
  auto blocks = [&](long long total) { long long b = (total + threads - 1) / threads; const long long cap = (long long)c.sm_count * 32; return (unsigned)(b < cap ? (b > 0 ? b : 1) : cap); };
  // node i's output for all n inputs: [n][H][W][C] at arena + act_off * n; the input node is the caller's buffer and the
  // output node (the softmax layer's input) goes straight to d_logits
  auto act = [&](int i) -> int8_t * {
    if (i == 0) return const_cast<int8_t *>(d_in);
    if (i == m->out_node) return d_logits;
    return m->d_arena + m->N[i].act_off * (long long)n;
  };
  // tensor-core path; word loads when the input channels (dense: the flattened input) are a multiple of 4 and the buffer is 4-byte aligned
  auto run_mma = [&](const GenNode &L0, const int8_t *src, int8_t *dst, bool dense) -> bool {
    GenNode L = L0;
    if (dense) { L.aC = L0.aH * L0.aW * L0.aC; L.aH = L.aW = 1; L.kh = L.kw = L.sh = L.sw = 1; L.ph = L.pw = 0; }
    if (m->naive) return false;
    const int K = L.kh * L.kw * L.aC;
    const bool bytes = (L.aC & 3) || (reinterpret_cast<uintptr_t>(src) & 3);
    const size_t smem = (size_t)(bytes ? (K + 31) & ~31 : ((K >> 2) + 7) & ~7) * sizeof(int2);
    if (smem > 48 * 1024) return false;
    const long long n_rows = (long long)n * L.H * L.W, tiles = (n_rows + 31) / 32;
    long long ctas = (tiles + kGenMmaWarps - 1) / kGenMmaWarps;
    const long long cap = (long long)c.sm_count * 8;
    if (ctas > cap) ctas = cap;
    const int bounds = L.same || L.ph || L.pw;
    const int8_t *w = m->d_w + L.w_off, *b = m->d_w + L.b_off;
    const unsigned grid = (unsigned)ctas, thr = kGenMmaWarps * 32;
    if (bytes) {
      if (L.C <= 16) gen_conv_mma_kernel<2, true><<<grid, thr, smem, st>>>(src, dst, w, b, L, n_rows, K, bounds);
      else if (L.C <= 32) gen_conv_mma_kernel<4, true><<<grid, thr, smem, st>>>(src, dst, w, b, L, n_rows, K, bounds);
      else gen_conv_mma_kernel<8, true><<<grid, thr, smem, st>>>(src, dst, w, b, L, n_rows, K, bounds);
    } else {
      if (L.C <= 16) gen_conv_mma_kernel<2, false><<<grid, thr, smem, st>>>(src, dst, w, b, L, n_rows, K, bounds);
      else if (L.C <= 32) gen_conv_mma_kernel<4, false><<<grid, thr, smem, st>>>(src, dst, w, b, L, n_rows, K, bounds);
      else gen_conv_mma_kernel<8, false><<<grid, thr, smem, st>>>(src, dst, w, b, L, n_rows, K, bounds);
    }
    return true;
  };
  for (int i = 1; i < m->n_nodes; i++) {
    const GenNode &L = m->N[i];
    const long long total = (long long)n * L.H * L.W * L.C;
    int8_t *dst = act(i);
    switch (L.op) {
      case G_CONV:
        if (!run_mma(L, act(L.in[0]), dst, false))
          gen_conv_kernel<<<blocks(total), threads, 0, st>>>(act(L.in[0]), dst, m->d_w + L.w_off, m->d_w + L.b_off, L, (long long)n);
        break;
      case G_DENSE:
        if (!run_mma(L, act(L.in[0]), dst, true))
          gen_dense_kernel<<<blocks(total), threads, 0, st>>>(act(L.in[0]), dst, m->d_w + L.w_off, m->d_w + L.b_off, L, (long long)n);
        break;
      case G_MAXPOOL:
      case G_GMAXPOOL: gen_maxpool_kernel<<<blocks(total), threads, 0, st>>>(act(L.in[0]), dst, L, (long long)n); break;
      case G_CONCAT: {
        ConcatSrc s;
        for (int k = 0; k < 8; k++) s.p[k] = k < L.n_in ? act(L.in[k]) : nullptr;
        gen_concat_kernel<<<blocks(total), threads, 0, st>>>(s, dst, L, (long long)n);
        break;
      }
      case G_ADD: gen_add_kernel<<<blocks(total), threads, 0, st>>>(act(L.in[0]), act(L.in[1]), dst, L, (long long)n); break;
      case G_SOFTMAX: continue;         // below, in the context's softmax variant
      default: return ctx_fail(EDISON_E_MODEL, "generic NNoM executor: node %d has op %d", i, L.op);
    }
    count_launch();
  }
  if (d_softmax) {
    gen_softmax_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_logits, d_softmax, m->n_out, (long long)n, c.softmax_mode);
    count_launch();
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "generic NNoM executor: %s", cudaGetErrorString(e));
  return EDISON_OK;
}

static bool gen_dev_ptr(const void *p) {
  cudaPointerAttributes a;
  if (!p || cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

static int ceil_div(int a, int b) { return (a + b - 1) / b; }

// append a node; weights (if any) are copied into wbuf.  Returns 0 or an error string
static const char *add_node(edison_model *m, std::vector<int8_t> &wbuf, GenNode L, const int8_t *w, size_t nw, const int8_t *b, size_t nb) {
  if (m->n_nodes >= kMaxNodes) return "too many nodes";
  const int i = m->n_nodes;
  for (int k = 0; k < L.n_in; k++)
    if (L.in[k] < 0 || L.in[k] >= i) return "input index out of order";
  const GenNode *A = L.n_in ? &m->N[L.in[0]] : nullptr;
  if (L.op != G_INPUT && !A) return "node without input";
  if (A) { L.aH = A->H; L.aW = A->W; L.aC = A->C; }
  switch (L.op) {
    case G_INPUT: break;
    case G_CONV:
      if (L.kh < 1 || L.kw < 1 || L.sh < 1 || L.sw < 1 || L.rshift < 1 || L.rshift > 31 || L.lshift < 0 || L.lshift > 24) return "conv parameters out of range";
      L.H = L.same ? ceil_div(A->H, L.sh) : ceil_div(A->H - L.kh + 1, L.sh);
      L.W = L.same ? ceil_div(A->W, L.sw) : ceil_div(A->W - L.kw + 1, L.sw);
      L.C = L.cout;
      if (nw != (size_t)L.cout * L.kh * L.kw * A->C || nb != (size_t)L.cout) return "conv weights do not fit the input channels";
      break;
    case G_DENSE:
      if (L.rshift < 1 || L.rshift > 31 || L.lshift < 0 || L.lshift > 24) return "dense shifts out of range";
      L.H = L.W = 1; L.C = L.cout;
      if (nw != (size_t)L.cout * A->H * A->W * A->C || nb != (size_t)L.cout) return "dense weights do not fit the input";
      break;
    case G_MAXPOOL:
      if (L.kh < 1 || L.kw < 1 || L.sh < 1 || L.sw < 1) return "pool parameters out of range";
      L.H = L.same ? ceil_div(A->H, L.sh) : ceil_div(A->H - L.kh + 1, L.sh);
      L.W = L.same ? ceil_div(A->W, L.sw) : ceil_div(A->W - L.kw + 1, L.sw);
      L.C = A->C;
      break;
    case G_GMAXPOOL:
      L.kh = A->H; L.kw = A->W; L.sh = L.sw = 1; L.ph = L.pw = 0;
      L.H = L.W = 1; L.C = A->C;
      break;
    case G_CONCAT:
      if (L.n_in < 2) return "concat needs two inputs or more";
      for (int k = 1; k < L.n_in; k++) {
        const GenNode &B = m->N[L.in[k]];
        if (B.H != A->H || B.W != A->W || B.C != A->C) return "concat inputs differ in shape (not reproducible in the reference runtime)";
      }
      L.H = A->H; L.W = A->W; L.C = A->C * L.n_in;
      break;
    case G_ADD: {
      if (L.n_in != 2 || L.rshift < 1 || L.rshift > 15) return "add needs two inputs and a shift of 1..15";
      const GenNode &B = m->N[L.in[1]];
      if (B.H != A->H || B.W != A->W || B.C != A->C) return "add inputs differ in shape";
      L.H = A->H; L.W = A->W; L.C = A->C;
      break;
    }
    case G_SOFTMAX: L.H = A->H; L.W = A->W; L.C = A->C; break;
    default: return "unknown op";
  }
  if (L.H < 1 || L.W < 1 || L.C < 1) return "empty output";
  if (nw) {
    L.w_off = (long long)wbuf.size();
    wbuf.insert(wbuf.end(), w, w + nw);
    wbuf.resize((wbuf.size() + 3) & ~(size_t)3);
    L.b_off = (long long)wbuf.size();
    wbuf.insert(wbuf.end(), b, b + nb);
    wbuf.resize((wbuf.size() + 3) & ~(size_t)3);
  }
  L.act_size = ((long long)L.H * L.W * L.C + 3) & ~3LL;
  L.act_off = m->arena_per_input;
  if (L.op != G_INPUT && L.op != G_SOFTMAX) m->arena_per_input += L.act_size;
  m->N[i] = L;
  m->n_nodes = i + 1;
  return nullptr;
}

static GenNode blank_node(int op, int n_in) {
  GenNode L;
  memset(&L, 0, sizeof L);
  L.op = op; L.n_in = n_in;
  for (int k = 0; k < 8; k++) L.in[k] = -1;
  return L;
}

static const char *parse_v1(edison_model *m, std::vector<int8_t> &wbuf, const std::vector<unsigned char> &raw, uint32_t nl) {
  // the trailer holds the input shape: walk the layers once to find it
  size_t off = 16;
  for (uint32_t i = 0; i < nl; i++) {
    uint32_t h[9];
    if (off + 36 > raw.size()) return "truncated";
    memcpy(h, &raw[off], 36);
    // every factor is bounded by the blob size before it is multiplied in: a hostile header cannot wrap size_t
    size_t nw = 1;
    for (int k = 1; k <= 4; k++) {
      if (h[k] == 0 || h[k] > raw.size() || nw > raw.size() / h[k]) return "truncated (layer larger than the blob)";
      nw *= h[k];
    }
    if (nw + h[1] > raw.size() - off - 36) return "truncated";
    off += 36 + nw + h[1];
    off = (off + 3) & ~(size_t)3;
  }
  if (off + 24 > raw.size()) return "trailer missing";
  uint32_t s[3];
  memcpy(s, &raw[off + 12], 12);
  GenNode in = blank_node(G_INPUT, 0);
  in.H = (int)s[0]; in.W = (int)s[1]; in.C = (int)s[2];
  const char *err = add_node(m, wbuf, in, nullptr, 0, nullptr, 0);
  if (err) return err;
  off = 16;
  for (uint32_t i = 0; i < nl; i++) {
    uint32_t h[9];
    memcpy(h, &raw[off], 36); off += 36;
    const size_t nw = (size_t)h[1] * h[2] * h[3] * h[4];
    const int8_t *w = (const int8_t *)&raw[off], *b = w + nw;
    off = (off + nw + h[1] + 3) & ~(size_t)3;
    if (h[0] != 1 && h[0] != 2) return "layer kind unsupported";
    GenNode L = blank_node(h[0] == 1 ? G_CONV : G_DENSE, 1);
    L.in[0] = m->n_nodes - 1;
    L.cout = (int)h[1]; L.kh = (int)h[2]; L.kw = (int)h[3]; L.sh = L.sw = 1;
    L.rshift = (int)h[5]; L.lshift = (int)h[6]; L.relu = h[7] ? 1 : 0;
    const GenNode &A = m->N[L.in[0]];
    if (h[0] == 1 && ((int)h[4] != A.C || L.kh > A.H || L.kw > A.W)) return "conv layer does not fit its input";
    if ((err = add_node(m, wbuf, L, w, nw, b, h[1]))) return err;
    if (h[0] == 1 && h[8]) {                       // (pool x 1) max pool, stride (pool, 1), valid
      GenNode P = blank_node(G_MAXPOOL, 1);
      P.in[0] = m->n_nodes - 1;
      P.kh = P.sh = (int)h[8]; P.kw = P.sw = 1;
      if (m->N[P.in[0]].H < P.kh) return "layer pools away its output";
      if ((err = add_node(m, wbuf, P, nullptr, 0, nullptr, 0))) return err;
    }
  }
  return nullptr;
}

static const char *parse_v2(edison_model *m, std::vector<int8_t> &wbuf, const std::vector<unsigned char> &raw, uint32_t nn) {
  size_t off = 16;
  for (uint32_t i = 0; i < nn; i++) {
    uint32_t h[20];
    if (off + 80 > raw.size()) return "truncated";
    memcpy(h, &raw[off], 80); off += 80;
    if (h[1] > 8) return "too many inputs";
    GenNode L = blank_node((int)h[0], (int)h[1]);
    for (uint32_t k = 0; k < h[1]; k++) L.in[k] = (int)h[2 + k];
    const uint32_t *p = h + 10;
    size_t nw = 0, nb = 0;
    switch (L.op) {
      case G_INPUT: L.H = (int)p[0]; L.W = (int)p[1]; L.C = (int)p[2]; if (i) return "input node must come first"; break;
      case G_CONV:
        L.cout = (int)p[0]; L.kh = (int)p[1]; L.kw = (int)p[2]; L.sh = (int)p[3]; L.sw = (int)p[4];
        if (p[5]) { L.same = 1; L.ph = (L.kh - 1) / 2; L.pw = (L.kw - 1) / 2; }      // nnom_conv2d.c:77-82
        L.rshift = (int)p[6]; L.lshift = (int)p[7]; L.relu = p[8] ? 1 : 0;
        nw = (size_t)p[0] * p[1] * p[2] * p[9]; nb = p[0];
        break;
      case G_DENSE:
        L.cout = (int)p[0]; L.rshift = (int)p[2]; L.lshift = (int)p[3]; L.relu = p[4] ? 1 : 0;
        nw = (size_t)p[0] * p[1]; nb = p[0];
        break;
      case G_MAXPOOL:
        L.kh = (int)p[0]; L.kw = (int)p[1]; L.sh = (int)p[2]; L.sw = (int)p[3];
        if (p[4]) { L.same = 1; L.ph = (L.kh - 1) / 2; L.pw = (L.kw - 1) / 2; }      // nnom_maxpool.c:66-71
        break;
      case G_ADD: L.rshift = (int)p[0]; L.relu = p[1] ? 1 : 0; break;
      case G_SOFTMAX: if (i != nn - 1) return "softmax must be the last node"; break;
      default: break;
    }
    if (i == 0 && L.op != G_INPUT) return "input node must come first";
    if (nw > raw.size() || off + nw + nb > raw.size()) return "truncated";
    const int8_t *w = (const int8_t *)&raw[off];
    const char *err = add_node(m, wbuf, L, w, nw, w + nw, nb);
    if (err) return err;
    if (nw) off = (off + nw + nb + 3) & ~(size_t)3;
  }
  if (off + 12 > raw.size()) return "trailer missing";
  return nullptr;
}

}  // namespace edison

using namespace edison;

extern "C" {

int edison_model_load(const char *blob_path, edison_model **out) {
  DeviceScope device_scope_;
  if (!blob_path || !out) return ctx_fail(EDISON_E_ARG, "edison_model_load: null argument");
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  ctx_unlock();
  if (rc) return rc;
  FILE *f = fopen(blob_path, "rb");
  if (!f) return ctx_fail(EDISON_E_DATA, "cannot open %s", blob_path);
  std::vector<unsigned char> raw;
  unsigned char buf[65536];
  size_t got;
  while ((got = fread(buf, 1, sizeof buf, f)) > 0) raw.insert(raw.end(), buf, buf + got);
  fclose(f);
  if (raw.size() < 16 || memcmp(raw.data(), "EDNNQ7\0\0", 8) != 0) return ctx_fail(EDISON_E_DATA, "%s: not an NNoM q7 blob", blob_path);
  uint32_t ver, nl;
  memcpy(&ver, &raw[8], 4); memcpy(&nl, &raw[12], 4);
  if ((ver != 1 && ver != 2) || nl < 1 || nl > (ver == 1 ? 16u : (uint32_t)kMaxNodes))
    return ctx_fail(EDISON_E_MODEL, "%s: version %u / %u layers unsupported", blob_path, ver, nl);
  edison_model *m = new edison_model();
  memset(m, 0, sizeof *m);
  std::vector<int8_t> wbuf;
  const char *err = ver == 1 ? parse_v1(m, wbuf, raw, nl) : parse_v2(m, wbuf, raw, nl);
  if (!err && m->n_nodes < 2) err = "no layers";
  if (err) {
    const int at = m->n_nodes;
    delete m;
    return ctx_fail(strstr(err, "truncated") || strstr(err, "trailer") ? EDISON_E_DATA : EDISON_E_MODEL, "%s: node %d: %s", blob_path, at, err);
  }
  const GenNode &last = m->N[m->n_nodes - 1];
  m->out_node = last.op == G_SOFTMAX ? last.in[0] : m->n_nodes - 1;
  if (m->out_node < 1) { delete m; return ctx_fail(EDISON_E_MODEL, "%s: no output layer", blob_path); }
  for (int i = 1; i < m->n_nodes; i++)
    for (int k = 0; k < m->N[i].n_in; k++)
      if (m->N[i].in[k] == m->out_node && m->N[i].op != G_SOFTMAX) { delete m; return ctx_fail(EDISON_E_MODEL, "%s: the output node has other readers", blob_path); }
  m->in_h = m->N[0].H; m->in_w = m->N[0].W; m->in_c = m->N[0].C;
  m->in_size = m->in_h * m->in_w * m->in_c;
  m->n_out = m->N[m->out_node].H * m->N[m->out_node].W * m->N[m->out_node].C;
  if (m->arena_per_input < 4) m->arena_per_input = 4;
  cudaError_t e = cudaMalloc(&m->d_w, wbuf.size() + 16);
  if (e == cudaSuccess) e = cudaMemcpy(m->d_w, wbuf.data(), wbuf.size(), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { edison_model_free(m); return ctx_fail(EDISON_E_CUDA, "edison_model_load: %s", cudaGetErrorString(e)); }
  *out = m;
  return EDISON_OK;
}

void edison_model_free(edison_model *m) {
  DeviceScope device_scope_;
  if (!m) return;
  cudaDeviceSynchronize();
  void *bufs[] = {m->d_w, m->d_arena, m->d_in, m->d_lg, m->d_sm};
  for (void *b : bufs)
    if (b) cudaFree(b);
  delete m;
}

int edison_model_shape(const edison_model *m, int *in_h, int *in_w, int *in_c, int *n_out) {
  if (!m) return ctx_fail(EDISON_E_ARG, "edison_model_shape: null model");
  if (in_h) *in_h = m->in_h;
  if (in_w) *in_w = m->in_w;
  if (in_c) *in_c = m->in_c;
  if (n_out) *n_out = m->n_out;
  return EDISON_OK;
}

int edison_model_infer(edison_model *m, const int8_t *in, size_t n, int8_t *logits, int8_t *softmax) {
  DeviceScope device_scope_;
  if (!m || (n && !in)) return ctx_fail(EDISON_E_ARG, "edison_model_infer: null argument");
  if (!n) return EDISON_OK;
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  if (!rc) rc = model_reserve(m, n);
  if (!rc) {
    const bool dev = gen_dev_ptr(in);
    const int8_t *d_in = in;
    cudaError_t e = cudaSuccess;
    if (!dev) {
      e = cudaMemcpyAsync(m->d_in, in, n * (size_t)m->in_size, cudaMemcpyHostToDevice, c.stream);
      d_in = m->d_in;
    }
    if (e != cudaSuccess) rc = ctx_fail(EDISON_E_CUDA, "edison_model_infer upload: %s", cudaGetErrorString(e));
    int8_t *d_lg = dev && logits ? logits : m->d_lg;
    int8_t *d_sm = softmax ? (dev ? softmax : m->d_sm) : nullptr;
    if (!rc) rc = model_run(m, c, d_in, n, d_lg, d_sm, c.stream);
    if (!rc && !dev) {
      if (logits) e = cudaMemcpyAsync(logits, d_lg, n * (size_t)m->n_out, cudaMemcpyDeviceToHost, c.stream);
      if (e == cudaSuccess && softmax) e = cudaMemcpyAsync(softmax, d_sm, n * (size_t)m->n_out, cudaMemcpyDeviceToHost, c.stream);
    }
    if (!rc && e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
    if (!rc && e != cudaSuccess) rc = ctx_fail(EDISON_E_CUDA, "edison_model_infer: %s", cudaGetErrorString(e));
  }
  ctx_unlock();
  return rc;
}

// debug / measurement: 1 = run every node on the one-thread-per-output kernels (the round-1 executor), 0 = tensor-core nodes where they apply
int edison_model_set_naive(edison_model *m, int naive) {
  if (!m) return ctx_fail(EDISON_E_ARG, "edison_model_set_naive: null model");
  ctx_lock();
  m->naive = naive ? 1 : 0;
  ctx_unlock();
  return EDISON_OK;
}

}  // extern "C"
