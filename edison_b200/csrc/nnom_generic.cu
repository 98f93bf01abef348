// nnom_generic.cu -- generic executor for NNoM int8 models of the sequential conv/dense family.
//
// K2 (nnom_q7_kernel.cu) is specialised for the reference's kws model.  Any other model that nnom_utils.py generates from the
// same layer vocabulary -- Input, Conv2D (valid, stride 1), ReLU, MaxPool (p,1)/(p,1), Dense, Softmax -- is converted by
// tools/convert_nnom_weights.py into the same blob format and runs here layer by layer with activations in HBM: one thread
// per output element, dp4a over the channel axis when it is a multiple of 4.  Same arithmetic as K2 and as the reference
// (local_convolve_HWC_q7_nonsquare nnom_local.c:425-484, local_maxpool_q7_HWC :117-160, local_fully_connected_q7 :733-802,
// softmax variants as in nnom_q7_kernel.cu), so results are bit-exact; it is a correctness-first fallback, not a tuned kernel.
// Pinned against the upstream NNoM keyword_spotting example model (audio/edison/deploy/nnom/examples/keyword_spotting/mcu/
// weights.h: 63x12 input, 35 classes) run through the reference's own NNoM runtime built for the host.
#include <stdio.h>
#include <string.h>

#include <vector>

#include "edison_internal.h"
#include "../../include/edison_b200.h"

namespace edison {

struct GenLayer {
  int kind, cout, kh, kw, cin, rshift, lshift, relu, pool;
  int in_h, in_w, out_h, out_w;       // conv: out_h is after pooling
  long long w_off, b_off;             // into the device weight buffer
};

__device__ __forceinline__ int g_ssat8(int v) { return min(max(v, -128), 127); }
__device__ __forceinline__ int g_usat(int v, int bits) { return min(max(v, 0), (1 << bits) - 1); }

// conv (valid, stride 1) + optional ReLU + optional (pool x 1) max pool.  in: [n][H][W][C], out: [n][HP][WO][O]
__global__ void gen_conv_kernel(const int8_t *__restrict__ in, int8_t *__restrict__ out, const int8_t *__restrict__ w,
                                const int8_t *__restrict__ b, GenLayer L, long long n) {
  const int O = L.cout, C = L.cin, W = L.in_w, WO = L.out_w, HP = L.out_h, P = L.pool ? L.pool : 1;
  const long long total = n * HP * WO * O;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int o = (int)(idx % O);
    long long r = idx / O;
    const int x = (int)(r % WO); r /= WO;
    const int yp = (int)(r % HP);
    const long long clip = r / HP;
    const int8_t *src = in + clip * ((long long)L.in_h * W * C);
    const int8_t *wo = w + (long long)o * L.kh * L.kw * C;
    const int seed = (int)((unsigned)(int)b[o] << L.lshift) + (1 << (L.rshift - 1));
    int best = -129;
    for (int p = 0; p < P; p++) {
      const int y = yp * P + p;
      int acc = seed;
      for (int ky = 0; ky < L.kh; ky++) {
        const int8_t *row = src + ((long long)(y + ky) * W + x) * C;
        const int8_t *wr = wo + ky * L.kw * C;
        const int len = L.kw * C;            // contiguous in both operands (HWC activations, O,H,W,I weights)
        int i = 0;
        if ((C & 3) == 0) {
          for (; i < len; i += 4)
            acc = __dp4a(*reinterpret_cast<const int *>(row + i), *reinterpret_cast<const int *>(wr + i), acc);
        } else {
          for (; i < len; i++) acc += (int)row[i] * (int)wr[i];
        }
      }
      int v = g_ssat8(acc >> L.rshift);
      if (L.relu && v < 0) v = 0;
      best = max(best, v);
    }
    out[idx] = (int8_t)best;
  }
}

// dense: in [n][cin] (the flattened H,W,C output of the previous layer), w [cout][cin]
__global__ void gen_dense_kernel(const int8_t *__restrict__ in, int8_t *__restrict__ out, const int8_t *__restrict__ w,
                                 const int8_t *__restrict__ b, GenLayer L, long long n) {
  const long long total = n * L.cout;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int o = (int)(idx % L.cout);
    const long long clip = idx / L.cout;
    const int8_t *x = in + clip * L.cin, *wr = w + (long long)o * L.cin;
    int acc = (int)((unsigned)(int)b[o] << L.lshift) + (1 << (L.rshift - 1));
    int i = 0;
    if ((L.cin & 3) == 0)
      for (; i < L.cin; i += 4) acc = __dp4a(*reinterpret_cast<const int *>(x + i), *reinterpret_cast<const int *>(wr + i), acc);
    else
      for (; i < L.cin; i++) acc += (int)x[i] * (int)wr[i];
    int v = g_ssat8(acc >> L.rshift);
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
  int n_layers;
  edison::GenLayer L[16];
  int8_t *d_w;               // all weights and biases, 4-byte aligned per array
  int in_h, in_w, in_c, in_size, n_out;
  long long max_act;         // bytes per clip of the largest activation
  int8_t *d_act[2], *d_in, *d_lg, *d_sm;
  size_t cap;                // clips the scratch buffers hold
};

namespace edison {

static int model_reserve(edison_model *m, size_t n) {
  if (n <= m->cap) return EDISON_OK;
  int8_t **bufs[] = {&m->d_act[0], &m->d_act[1], &m->d_in, &m->d_lg, &m->d_sm};
  const size_t per[] = {(size_t)m->max_act, (size_t)m->max_act, (size_t)m->in_size, (size_t)m->n_out, (size_t)m->n_out};
  for (int i = 0; i < 5; i++) {
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
  const int8_t *cur = d_in;
  const int threads = 256;
  auto blocks = [&](long long total) { long long b = (total + threads - 1) / threads; const long long cap = (long long)c.sm_count * 32; return (unsigned)(b < cap ? (b > 0 ? b : 1) : cap); };
  for (int i = 0; i < m->n_layers; i++) {
    const GenLayer &L = m->L[i];
    int8_t *dst = (i == m->n_layers - 1) ? d_logits : m->d_act[i & 1];
    if (L.kind == 1) {
      const long long total = (long long)n * L.out_h * L.out_w * L.cout;
      gen_conv_kernel<<<blocks(total), threads, 0, st>>>(cur, dst, m->d_w + L.w_off, m->d_w + L.b_off, L, (long long)n);
    } else {
      gen_dense_kernel<<<blocks((long long)n * L.cout), threads, 0, st>>>(cur, dst, m->d_w + L.w_off, m->d_w + L.b_off, L, (long long)n);
    }
    count_launch();
    cur = dst;
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

}  // namespace edison

using namespace edison;

extern "C" {

int edison_model_load(const char *blob_path, edison_model **out) {
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
  if (ver != 1 || nl < 1 || nl > 16) return ctx_fail(EDISON_E_MODEL, "%s: version %u / %u layers unsupported", blob_path, ver, nl);
  edison_model *m = new edison_model();
  memset(m, 0, sizeof *m);
  m->n_layers = (int)nl;
  std::vector<int8_t> wbuf;
  size_t off = 16;
  for (uint32_t i = 0; i < nl; i++) {
    uint32_t h[9];
    if (off + 36 > raw.size()) { delete m; return ctx_fail(EDISON_E_DATA, "%s: truncated", blob_path); }
    memcpy(h, &raw[off], 36); off += 36;
    GenLayer &L = m->L[i];
    L.kind = h[0]; L.cout = h[1]; L.kh = h[2]; L.kw = h[3]; L.cin = h[4]; L.rshift = h[5]; L.lshift = h[6]; L.relu = h[7]; L.pool = h[8];
    const size_t nw = (size_t)L.cout * L.kh * L.kw * L.cin;
    if ((L.kind != 1 && L.kind != 2) || L.rshift < 1 || L.rshift > 31 || L.lshift > 24 || off + nw + L.cout > raw.size()) {
      delete m; return ctx_fail(EDISON_E_MODEL, "%s: layer %u unsupported or truncated", blob_path, i);
    }
    L.w_off = (long long)wbuf.size();
    wbuf.insert(wbuf.end(), (int8_t *)&raw[off], (int8_t *)&raw[off] + nw);
    wbuf.resize((wbuf.size() + 3) & ~(size_t)3);
    off += nw;
    L.b_off = (long long)wbuf.size();
    wbuf.insert(wbuf.end(), (int8_t *)&raw[off], (int8_t *)&raw[off] + L.cout);
    wbuf.resize((wbuf.size() + 3) & ~(size_t)3);
    off += L.cout;
    off = (off + 3) & ~(size_t)3;
  }
  if (off + 24 > raw.size()) { delete m; return ctx_fail(EDISON_E_DATA, "%s: trailer missing", blob_path); }
  uint32_t s[3];
  memcpy(s, &raw[off + 12], 12);
  m->in_h = s[0]; m->in_w = s[1]; m->in_c = s[2];
  m->in_size = m->in_h * m->in_w * m->in_c;
  // shape propagation and checks
  int H = m->in_h, W = m->in_w, C = m->in_c;
  m->max_act = 4;
  for (int i = 0; i < m->n_layers; i++) {
    GenLayer &L = m->L[i];
    if (L.kind == 1) {
      if (L.cin != C || L.kh > H || L.kw > W) { delete m; return ctx_fail(EDISON_E_MODEL, "%s: conv layer %d does not fit its input (%dx%dx%d)", blob_path, i, H, W, C); }
      L.in_h = H; L.in_w = W;
      const int ho = H - L.kh + 1;
      L.out_w = W - L.kw + 1;
      L.out_h = L.pool ? (ho - L.pool) / L.pool + 1 : ho;
      if (L.out_h < 1) { delete m; return ctx_fail(EDISON_E_MODEL, "%s: layer %d pools away its output", blob_path, i); }
      H = L.out_h; W = L.out_w; C = L.cout;
    } else {
      if (L.cin != H * W * C) { delete m; return ctx_fail(EDISON_E_MODEL, "%s: dense layer %d expects %d inputs, gets %d", blob_path, i, L.cin, H * W * C); }
      L.in_h = 1; L.in_w = 1; L.out_h = 1; L.out_w = 1;
      H = 1; W = 1; C = L.cout;
    }
    const long long act = (long long)H * W * C;
    if (act > m->max_act) m->max_act = act;
  }
  m->max_act = (m->max_act + 3) & ~3LL;
  m->n_out = H * W * C;
  cudaError_t e = cudaMalloc(&m->d_w, wbuf.size() + 16);
  if (e == cudaSuccess) e = cudaMemcpy(m->d_w, wbuf.data(), wbuf.size(), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { edison_model_free(m); return ctx_fail(EDISON_E_CUDA, "edison_model_load: %s", cudaGetErrorString(e)); }
  *out = m;
  return EDISON_OK;
}

void edison_model_free(edison_model *m) {
  if (!m) return;
  cudaDeviceSynchronize();
  void *bufs[] = {m->d_w, m->d_act[0], m->d_act[1], m->d_in, m->d_lg, m->d_sm};
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

}  // extern "C"
