/* nnom_graph.c -- TEST INFRASTRUCTURE: CPU restatement of the reference's NNoM runtime (local backend) for models whose
 * graph is more than the reference model's conv/dense chain.  Input: the version-2 blob of tools/convert_nnom_graph.py.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it.
 *
 * Restated functions (firmware/src/ai/nnom/src, the runtime the reference vendors):
 *   conv      layers/nnom_conv2d.c:77-82 (SAME pad = (k-1)/2), :101-112 (output size), backends/nnom_local.c:425-483
 *   ReLU tail backends/nnom_local.c:1075-1085 via layers/nnom_activation.c
 *   max pool  layers/nnom_maxpool.c:66-77 (pad), :92-104 (output size), backends/nnom_local.c:117-159
 *   global max pool  layers/nnom_global_pool.c:93-96 (kernel = whole map, stride 1, no pad) -> the same max pool
 *   concat    layers/nnom_concat.c:186-213 (HWC: for every pixel, the channel vector of each input in hook order;
 *             every input is sized like the first one, so only equal shapes are meaningful -- the converter refuses others)
 *   add       layers/nnom_matrix.c add_run -> backends/nnom_local.c:1109-1126 (rounding shift, saturate to 8 bits)
 *   dense     layers/nnom_dense.c:92-119 -> backends/nnom_local.c:804-921 (weights de-interleaved by the converter)
 *   softmax   edison_oracle.c eo_softmax_q7 (nnom_local.c:948-1003 and the CMSIS-NN variants)
 * Pinned against the reference runtime itself on two upstream example models that the reference tree carries
 * (tests/golden/nnom_graphs.npz, made by tests/golden/make_golden_nnom_graphs.py from oracle/_ref/libref_nnom_ex_*.so). */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "edison_oracle.h"

enum { G_INPUT, G_CONV, G_DENSE, G_MAXPOOL, G_CONCAT, G_SOFTMAX, G_GMAXPOOL, G_ADD };
#define G_MAX_NODES 64

typedef struct {
  uint32_t op, n_in, in[8], p[10];
  const int8_t *w, *b;
  int H, W, C;
  int8_t *out;
} g_node;

static g_node nodes[G_MAX_NODES];
static uint32_t n_nodes;
static uint8_t *gblob;
static int out_node = -1, sm_node = -1;

static inline int32_t g_ssat8(int32_t v) { return v > 127 ? 127 : (v < -128 ? -128 : v); }
static int ceil_div(int a, int b) { return (a + b - 1) / b; }

static void graph_free(void)
{
  for (uint32_t i = 0; i < n_nodes; i++) { free(nodes[i].out); nodes[i].out = 0; }
  free(gblob); gblob = 0; n_nodes = 0; out_node = sm_node = -1;
}

int eo_graph_load(const char *path)
{
  FILE *fh = fopen(path, "rb");
  if (!fh) return -1;
  fseek(fh, 0, SEEK_END);
  long sz = ftell(fh);
  fseek(fh, 0, SEEK_SET);
  graph_free();
  gblob = (uint8_t *)malloc((size_t)sz);
  if (fread(gblob, 1, (size_t)sz, fh) != (size_t)sz) { fclose(fh); return -2; }
  fclose(fh);
  if (sz < 16 || memcmp(gblob, "EDNNQ7\0\0", 8)) return -3;
  uint32_t ver, nn;
  memcpy(&ver, gblob + 8, 4); memcpy(&nn, gblob + 12, 4);
  if (ver != 2 || nn < 2 || nn > G_MAX_NODES) return -4;
  const uint8_t *p = gblob + 16, *end = gblob + sz;
  for (uint32_t i = 0; i < nn; i++) {
    g_node *N = &nodes[i];
    if (p + 80 > end) return -5;
    memcpy(&N->op, p, 4); memcpy(&N->n_in, p + 4, 4); memcpy(N->in, p + 8, 32); memcpy(N->p, p + 40, 40);
    p += 80;
    if (N->n_in > 8) return -6;
    for (uint32_t k = 0; k < N->n_in; k++) if (N->in[k] >= i) return -6;       /* topological order */
    const g_node *A = N->n_in ? &nodes[N->in[0]] : 0;
    size_t nw = 0, nb = 0;
    switch (N->op) {
    case G_INPUT: N->H = (int)N->p[0]; N->W = (int)N->p[1]; N->C = (int)N->p[2]; break;
    case G_CONV: {
      const int kh = (int)N->p[1], kw = (int)N->p[2], sh = (int)N->p[3], sw = (int)N->p[4], same = (int)N->p[5];
      if (!A || (int)N->p[9] != A->C || !sh || !sw || N->p[6] < 1 || N->p[6] > 31 || N->p[7] > 24) return -7;
      N->H = same ? ceil_div(A->H, sh) : ceil_div(A->H - kh + 1, sh);
      N->W = same ? ceil_div(A->W, sw) : ceil_div(A->W - kw + 1, sw);
      N->C = (int)N->p[0];
      nw = (size_t)N->p[0] * kh * kw * A->C; nb = N->p[0];
      break; }
    case G_DENSE:
      if (!A || (int)N->p[1] != A->H * A->W * A->C || N->p[2] < 1 || N->p[2] > 31 || N->p[3] > 24) return -7;
      N->H = N->W = 1; N->C = (int)N->p[0];
      nw = (size_t)N->p[0] * N->p[1]; nb = N->p[0];
      break;
    case G_MAXPOOL: {
      const int kh = (int)N->p[0], kw = (int)N->p[1], sh = (int)N->p[2], sw = (int)N->p[3], same = (int)N->p[4];
      if (!A || !sh || !sw) return -7;
      N->H = same ? ceil_div(A->H, sh) : ceil_div(A->H - kh + 1, sh);
      N->W = same ? ceil_div(A->W, sw) : ceil_div(A->W - kw + 1, sw);
      N->C = A->C;
      break; }
    case G_GMAXPOOL: if (!A) return -7; N->H = N->W = 1; N->C = A->C; break;
    case G_CONCAT:
      if (N->n_in < 2) return -7;
      for (uint32_t k = 1; k < N->n_in; k++) {
        const g_node *B = &nodes[N->in[k]];
        if (B->H != A->H || B->W != A->W || B->C != A->C) return -8;
      }
      N->H = A->H; N->W = A->W; N->C = A->C * (int)N->n_in;
      break;
    case G_ADD: {
      if (N->n_in != 2 || N->p[0] < 1 || N->p[0] > 15) return -7;
      const g_node *B = &nodes[N->in[1]];
      if (B->H != A->H || B->W != A->W || B->C != A->C) return -8;
      N->H = A->H; N->W = A->W; N->C = A->C;
      break; }
    case G_SOFTMAX: if (!A || i != nn - 1) return -7; N->H = A->H; N->W = A->W; N->C = A->C; break;
    default: return -9;
    }
    if (N->H < 1 || N->W < 1 || N->C < 1) return -10;
    if (nw) {
      if (p + nw + nb > end) return -5;
      N->w = (const int8_t *)p; p += nw;
      N->b = (const int8_t *)p; p += nb;
      p += (size_t)(-(p - gblob)) & 3;
    }
    N->out = (int8_t *)malloc((size_t)N->H * N->W * N->C);
    n_nodes = i + 1;
  }
  if (p + 12 > end) return -5;
  sm_node = nodes[nn - 1].op == G_SOFTMAX ? (int)nn - 1 : -1;
  out_node = sm_node >= 0 ? (int)nodes[nn - 1].in[0] : (int)nn - 1;
  return 0;
}

int eo_graph_in_size(void) { return n_nodes ? nodes[0].H * nodes[0].W * nodes[0].C : 0; }
int eo_graph_out_size(void) { return out_node >= 0 ? nodes[out_node].H * nodes[out_node].W * nodes[out_node].C : 0; }
int eo_graph_n_nodes(void) { return (int)n_nodes; }
int eo_graph_node_size(int i) { return i >= 0 && (uint32_t)i < n_nodes ? nodes[i].H * nodes[i].W * nodes[i].C : -1; }
const int8_t *eo_graph_node_output(int i) { return i >= 0 && (uint32_t)i < n_nodes ? nodes[i].out : 0; }

/* local_convolve_HWC_q7_nonsquare, nnom_local.c:425-483: taps that fall outside the image are skipped */
static void g_conv(const g_node *N, const g_node *A)
{
  const int O = N->C, C = A->C, kh = (int)N->p[1], kw = (int)N->p[2], sh = (int)N->p[3], sw = (int)N->p[4];
  const int ph = N->p[5] ? (kh - 1) / 2 : 0, pw = N->p[5] ? (kw - 1) / 2 : 0;
  const int rs = (int)N->p[6], ls = (int)N->p[7], relu = (int)N->p[8];
  for (int o = 0; o < O; o++)
    for (int j = 0; j < N->H; j++)
      for (int k = 0; k < N->W; k++) {
        int32_t acc = (int32_t)((uint32_t)(int32_t)N->b[o] << ls) + (1 << (rs - 1));
        for (int m = 0; m < kh; m++)
          for (int n = 0; n < kw; n++) {
            const int r = sh * j + m - ph, c = sw * k + n - pw;
            if (r < 0 || c < 0 || r >= A->H || c >= A->W) continue;
            const int8_t *x = A->out + (r * A->W + c) * C, *w = N->w + ((o * kh + m) * kw + n) * C;
            for (int l = 0; l < C; l++) acc += x[l] * w[l];
          }
        int32_t v = g_ssat8(acc >> rs);
        if (relu && v < 0) v = 0;
        N->out[(j * N->W + k) * O + o] = (int8_t)v;
      }
}

/* local_maxpool_q7_HWC, nnom_local.c:117-159: starts from -129, stores the int as q7 */
static void g_maxpool(const g_node *N, const g_node *A, int kh, int kw, int sh, int sw, int ph, int pw)
{
  const int C = A->C;
  for (int c = 0; c < C; c++)
    for (int y = 0; y < N->H; y++)
      for (int x = 0; x < N->W; x++) {
        int mx = -129;
        for (int ky = y * sh - ph; ky < y * sh - ph + kh; ky++)
          for (int kx = x * sw - pw; kx < x * sw - pw + kw; kx++)
            if (ky >= 0 && kx >= 0 && ky < A->H && kx < A->W && A->out[c + C * (kx + ky * A->W)] > mx)
              mx = A->out[c + C * (kx + ky * A->W)];
        N->out[c + C * (x + y * N->W)] = (int8_t)mx;
      }
}

static int g_run(const int8_t *in)
{
  memcpy(nodes[0].out, in, (size_t)eo_graph_in_size());
  for (uint32_t i = 1; i < n_nodes; i++) {
    g_node *N = &nodes[i];
    const g_node *A = &nodes[N->in[0]];
    switch (N->op) {
    case G_CONV: g_conv(N, A); break;
    case G_DENSE: {
      const int n_in = (int)N->p[1], rs = (int)N->p[2], ls = (int)N->p[3], relu = (int)N->p[4];
      for (int r = 0; r < N->C; r++) {
        int32_t acc = (int32_t)((uint32_t)(int32_t)N->b[r] << ls) + (1 << (rs - 1));
        for (int j = 0; j < n_in; j++) acc += A->out[j] * N->w[r * n_in + j];
        int32_t v = g_ssat8(acc >> rs);
        if (relu && v < 0) v = 0;
        N->out[r] = (int8_t)v;
      }
      break; }
    case G_MAXPOOL: {
      const int kh = (int)N->p[0], kw = (int)N->p[1], same = (int)N->p[4];
      g_maxpool(N, A, kh, kw, (int)N->p[2], (int)N->p[3], same ? (kh - 1) / 2 : 0, same ? (kw - 1) / 2 : 0);
      break; }
    case G_GMAXPOOL: g_maxpool(N, A, A->H, A->W, 1, 1, 0, 0); break;
    case G_CONCAT: {
      int8_t *o = N->out;
      for (int px = 0; px < A->H * A->W; px++)
        for (uint32_t k = 0; k < N->n_in; k++) { memcpy(o, nodes[N->in[k]].out + px * A->C, (size_t)A->C); o += A->C; }
      break; }
    case G_ADD: {
      const g_node *B = &nodes[N->in[1]];
      const int rs = (int)N->p[0], relu = (int)N->p[1], sz = A->H * A->W * A->C;
      for (int k = 0; k < sz; k++) {
        int32_t v = g_ssat8(((int32_t)A->out[k] + B->out[k] + (1 << (rs - 1))) >> rs);
        if (relu && v < 0) v = 0;
        N->out[k] = (int8_t)v;
      }
      break; }
    case G_SOFTMAX: break;                       /* applied by the caller in the requested variant */
    default: return -1;
    }
  }
  return 0;
}

/* in [n][in_size] -> logits [n][out_size] (the softmax layer's input) and softmax [n][out_size] in variant `mode` */
int eo_graph_forward(const int8_t *in, size_t n, int8_t *logits, int8_t *softmax, int mode)
{
  if (!n_nodes) return -1;
  const size_t ni = (size_t)eo_graph_in_size(), no = (size_t)eo_graph_out_size();
  for (size_t i = 0; i < n; i++) {
    int r = g_run(in + i * ni);
    if (r) return r;
    if (logits) memcpy(logits + i * no, nodes[out_node].out, no);
    if (softmax) eo_softmax_q7(nodes[out_node].out, (int)no, softmax + i * no, mode);
  }
  return 0;
}
