// edison_api.cu -- the C ABI of libedison_b200.so (include/edison_b200.h), sections 0, 2, 3, 4.
//
// Host-side plumbing only: context, tables and model upload, pointer classification, and a
// two-slot chunk pipeline (H2D copy | kernels | D2H copy on alternating streams) for calls that pass
// host buffers.  All arithmetic happens in mfcc_q15_kernel.cu / nnom_q7_kernel.cu; there is no CPU
// implementation to fall back to, so every entry point fails loudly when CUDA is not usable.
#include <dlfcn.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <string>

#include "edison_internal.h"
#include <nvtx3/nvToolsExt.h>
#include "../../include/edison_b200.h"

namespace edison {

static std::atomic<unsigned long long> g_launches{0};
void count_launch(void) { g_launches.fetch_add(1, std::memory_order_relaxed); }

static thread_local char t_err[512] = "";
static int fail(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof t_err, fmt, ap);
  va_end(ap);
  return code;
}
#define CU(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) return fail(EDISON_E_CUDA, "%s: %s", #call, cudaGetErrorString(e_));    \
  } while (0)

struct Slot {
  cudaStream_t stream = nullptr;
  void *pcm = nullptr; size_t pcm_cap = 0;
  void *mfcc = nullptr; size_t mfcc_cap = 0;
  void *net_in = nullptr; size_t net_in_cap = 0;
  void *logits = nullptr; size_t logits_cap = 0;
  void *softmax = nullptr; size_t softmax_cap = 0;
};

struct Context {
  bool ready = false;
  int device = 0;
  int sm_count = 0;
  int softmax_mode = EDISON_SOFTMAX_CMSIS_REF;
  size_t chunk_clips = 8192;
  MfccTables *d_tables = nullptr;
  NnomModel *d_model = nullptr;
  int in_scale = 1, in_min = -128, in_max = 127;
  Slot slot[2];
  std::string data_dir;
  std::string keywords[32];
  int n_keywords = 0;
};
static Context g;
static std::mutex g_mu;

static int ensure(void **p, size_t *cap, size_t need) {
  if (need <= *cap) return EDISON_OK;
  if (*p) { cudaFree(*p); *p = nullptr; *cap = 0; }
  size_t want = need + need / 8;
  cudaError_t e = cudaMalloc(p, want);
  if (e != cudaSuccess) {
    cudaGetLastError();
    e = cudaMalloc(p, need);
    want = need;
  }
  if (e != cudaSuccess) return fail(EDISON_E_NOMEM, "cudaMalloc(%zu): %s", need, cudaGetErrorString(e));
  *cap = want;
  return EDISON_OK;
}

static std::string default_data_dir() {
  const char *env = getenv("EDISON_B200_DATA");
  if (env && *env) return env;
  Dl_info info;
  if (dladdr((void *)&count_launch, &info) && info.dli_fname) {
    std::string p = info.dli_fname;
    size_t s = p.rfind('/');
    std::string dir = s == std::string::npos ? "." : p.substr(0, s);
    return dir + "/data";
  }
  return "edison_b200/data";
}

static void load_keywords(const std::string &path) {
  g.n_keywords = 0;
  FILE *fh = fopen(path.c_str(), "r");
  if (!fh) return;
  std::string cur;
  bool in = false;
  int ch;
  while ((ch = fgetc(fh)) != EOF && g.n_keywords < 32) {
    if (ch == '"') {
      if (in) g.keywords[g.n_keywords++] = cur;
      cur.clear();
      in = !in;
    } else if (in) {
      cur.push_back((char)ch);
    }
  }
  fclose(fh);
}

static void release_device_state(void) {
  for (int i = 0; i < 2; i++) {
    Slot &s = g.slot[i];
    void *bufs[] = {s.pcm, s.mfcc, s.net_in, s.logits, s.softmax};
    for (void *b : bufs)
      if (b) cudaFree(b);
    if (s.stream) cudaStreamDestroy(s.stream);
    s = Slot();
  }
  if (g.d_tables) cudaFree(g.d_tables);
  nnom_free(g.d_model);
  g.d_tables = nullptr;
  g.d_model = nullptr;
}

// The firmware build is -DARM_MATH_DSP (firmware/Makefile:30) with NNOM_USING_CMSIS_NN (nnom_port.h:42): arm_softmax_q7 runs its
// SIMD branch (8-bit wrapping __SSUB8 / __SADD8), so that is what the drop-in symbols compute unless a cfg says otherwise.
static const int kDefaultSoftmax = EDISON_SOFTMAX_CMSIS_SIMD;

static int do_init_locked(const edison_cfg *cfg) {
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0) {
    cudaGetLastError();
    return fail(EDISON_E_NOGPU, "no usable CUDA device (%s); libedison_b200 has no CPU path",
                e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  }
  if (cfg && cfg->device >= 0) CU(cudaSetDevice(cfg->device));
  CU(cudaGetDevice(&g.device));
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, g.device));
  if (prop.major < 10)
    return fail(EDISON_E_NOGPU, "device %d is sm_%d%d; this library is built for sm_100a only", g.device, prop.major, prop.minor);
  g.sm_count = prop.multiProcessorCount;
  g.softmax_mode = cfg ? cfg->softmax_mode : kDefaultSoftmax;
  if (g.softmax_mode < 0 || g.softmax_mode > 2) return fail(EDISON_E_ARG, "softmax_mode %d", g.softmax_mode);
  g.chunk_clips = (cfg && cfg->chunk_clips) ? cfg->chunk_clips : 8192;
  g.data_dir = (cfg && cfg->data_dir) ? std::string(cfg->data_dir) : default_data_dir();

  char err[256] = "";
  NnomModel *dm = nullptr;
  int r = nnom_load((g.data_dir + "/kws_nnom_q7.bin").c_str(), &dm, &g.in_scale, &g.in_min, &g.in_max, err, sizeof err);
  if (r) return fail(r, "%s", err);
  g.d_model = dm;
  MfccTables *ht = (MfccTables *)malloc(sizeof(MfccTables));
  if (!ht) return fail(EDISON_E_NOMEM, "malloc");
  r = build_mfcc_tables(ht, (g.data_dir + "/mel_q15.bin").c_str(), err, sizeof err);
  if (r) { free(ht); return fail(r, "%s", err); }
  ht->in_scale = g.in_scale; ht->in_min = g.in_min; ht->in_max = g.in_max;
  cudaError_t ce = cudaMalloc(&g.d_tables, sizeof(MfccTables));
  if (ce == cudaSuccess) ce = cudaMemcpy(g.d_tables, ht, sizeof(MfccTables), cudaMemcpyHostToDevice);
  free(ht);
  if (ce != cudaSuccess) return fail(EDISON_E_CUDA, "table upload: %s", cudaGetErrorString(ce));
  load_keywords(g.data_dir + "/keywords.txt");
  CU(mfcc_q15_configure());
  CU(nnom_configure());
  for (int i = 0; i < 2; i++) CU(cudaStreamCreateWithFlags(&g.slot[i].stream, cudaStreamNonBlocking));
  return EDISON_OK;
}

static int do_init(const edison_cfg *cfg) {
  if (g.ready) {
    // one context per process: a second edison_init may repeat the first one's choices, not change them
    if (cfg) {
      if (cfg->device >= 0 && cfg->device != g.device)
        return fail(EDISON_E_ARG, "edison_init: already bound to device %d (asked for %d); edison_shutdown() first", g.device, cfg->device);
      if (cfg->softmax_mode != g.softmax_mode)
        return fail(EDISON_E_ARG, "edison_init: already initialised with softmax_mode %d (asked for %d); use edison_set_softmax_mode()",
                    g.softmax_mode, cfg->softmax_mode);
      if (cfg->data_dir && g.data_dir != cfg->data_dir)
        return fail(EDISON_E_ARG, "edison_init: already initialised from %s", g.data_dir.c_str());
    }
    return EDISON_OK;
  }
  int prev = -1;
  cudaGetDevice(&prev);
  cudaGetLastError();
  const int r = do_init_locked(cfg);
  if (r) {   // nothing half-built survives a failed init: a retry starts from scratch
    release_device_state();
    cudaGetLastError();
  } else {
    g.ready = true;
  }
  if (prev >= 0 && prev != g.device) cudaSetDevice(prev);   // the caller's current device is the caller's business
  return r;
}

int ctx_device(void) { return g.ready ? g.device : -1; }

TraceRange::TraceRange(const char *label) { nvtxRangePushA(label); }
TraceRange::~TraceRange() { nvtxRangePop(); }

// 0 = null, 1 = host (pageable or pinned), 2 = device / managed
static int ptr_kind(const void *p) {
  if (!p) return 0;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 1; }
  return (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) ? 2 : 1;
}
// every non-null pointer of a call must live on the same side as `ref`: a kernel storing to a pageable host address would be a
// sticky illegal-address fault for the whole process
static int same_side(const void *ref, const void *a, const void *b = nullptr, const void *c = nullptr, const void *d = nullptr) {
  const int k = ptr_kind(ref);
  const void *others[] = {a, b, c, d};
  for (const void *o : others)
    if (o && ptr_kind(o) != k)
      return fail(EDISON_E_ARG, "mixed host and device pointers in one call (input is a %s pointer)", k == 2 ? "device" : "host");
  return EDISON_OK;
}

static bool is_device_ptr(const void *p) {
  if (!p) return false;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// ---- device-resident cores -------------------------------------------------------------------------
static int mfcc_dev(const int16_t *d_pcm, size_t n_clips, size_t fpc, size_t stride, int16_t *d_mfcc, int8_t *d_net_in,
                    cudaStream_t st) {
  if (!n_clips || !fpc) return EDISON_OK;
  if (!d_pcm || (!d_mfcc && !d_net_in)) return fail(EDISON_E_ARG, "null pointer");
  if (fpc > 0x7fffffff) return fail(EDISON_E_ARG, "frames_per_clip too large");
  if (reinterpret_cast<uintptr_t>(d_pcm) & 1u) return fail(EDISON_E_ARG, "pcm must be 2-byte aligned");
  if (d_mfcc && (reinterpret_cast<uintptr_t>(d_mfcc) & 15u)) return fail(EDISON_E_ARG, "d_mfcc must be 16-byte aligned");
  MfccLaunch L;
  L.pcm = d_pcm; L.n_clips = (long long)n_clips; L.frames_per_clip = (int)fpc; L.clip_stride = (long long)stride;
  L.mfcc = d_mfcc; L.net_in = d_net_in; L.net_in_stride = EDISON_N_COEFF; L.tables = g.d_tables;
  TraceRange r("audioCalcMFCCs");   // audioprocessing.c:123
  CU(launch_mfcc_q15(L, g.sm_count, st));
  return EDISON_OK;
}

static int infer_dev(const int8_t *d_in, size_t n, int8_t *d_logits, int8_t *d_softmax, int32_t *d_counts, cudaStream_t st) {
  if (!n) return EDISON_OK;
  if (!d_in || (!d_logits && !d_softmax && !d_counts)) return fail(EDISON_E_ARG, "null pointer");
  K2Launch k{};
  k.net_in = d_in; k.n = (long long)n; k.logits = d_logits; k.softmax = d_softmax; k.softmax_mode = g.softmax_mode; k.class_counts = d_counts;
  TraceRange r("inference");        // app.c:207
  CU(launch_nnom_q7(g.d_model, k, g.sm_count, st));
  return EDISON_OK;
}

static int sync_slots(int rc) {
  for (int i = 0; i < 2; i++) {
    cudaError_t e = cudaStreamSynchronize(g.slot[i].stream);
    if (e != cudaSuccess && rc == EDISON_OK) rc = fail(EDISON_E_CUDA, "stream sync: %s", cudaGetErrorString(e));
  }
  return rc;
}

// ---- host-pointer pipeline ---------------------------------------------------------------------------
// One call = a sequence of chunks of clips; chunk k runs on slot k & 1, so the H2D copy of chunk k+1
// overlaps the kernels and the D2H copy of chunk k when the caller's buffers are pinned.
static int kws_host_chunks(const int16_t *pcm, size_t n_clips, size_t fpc, size_t stride, int16_t *mfcc, int8_t *net_in,
                           int8_t *logits, int8_t *softmax) {
  const bool want_net = logits || softmax;
  const size_t span = fpc * EDISON_FRAME_LEN;  // samples read from each clip
  const size_t target_bytes = (size_t)256 << 20;
  size_t chunk = g.chunk_clips;
  const size_t unit = (stride > span ? stride : span) * 2;
  if (chunk * unit > target_bytes) chunk = target_bytes / unit ? target_bytes / unit : 1;
  if (fpc == 1) chunk = target_bytes / unit ? target_bytes / unit : 1;  // flat frame batches: size by bytes only
  size_t k = 0;
  for (size_t c0 = 0; c0 < n_clips; c0 += chunk, k++) {
    const size_t nc = n_clips - c0 < chunk ? n_clips - c0 : chunk;
    Slot &s = g.slot[k & 1];
    const size_t in_samples = (nc - 1) * stride + span;
    const size_t nf = nc * fpc;
    int rc;
    if ((rc = ensure(&s.pcm, &s.pcm_cap, in_samples * 2))) return rc;
    if (mfcc && (rc = ensure(&s.mfcc, &s.mfcc_cap, nf * 64))) return rc;
    if ((net_in || want_net) && (rc = ensure(&s.net_in, &s.net_in_cap, nf * EDISON_N_COEFF))) return rc;
    if (logits && (rc = ensure(&s.logits, &s.logits_cap, nc * EDISON_N_CLASSES))) return rc;
    if (softmax && (rc = ensure(&s.softmax, &s.softmax_cap, nc * EDISON_N_CLASSES))) return rc;
    TraceRange chunk_range("appHifMfccAndInference");   // app.c:177; its events: "receive and MFCC" :199, "inference" :207
    CU(cudaMemcpyAsync(s.pcm, pcm + c0 * stride, in_samples * 2, cudaMemcpyHostToDevice, s.stream));
    int8_t *d_net = (net_in || want_net) ? (int8_t *)s.net_in : nullptr;
    if ((rc = mfcc_dev((const int16_t *)s.pcm, nc, fpc, stride, mfcc ? (int16_t *)s.mfcc : nullptr, d_net, s.stream))) return rc;
    if (want_net &&
        (rc = infer_dev(d_net, nc, logits ? (int8_t *)s.logits : nullptr, softmax ? (int8_t *)s.softmax : nullptr, nullptr, s.stream)))
      return rc;
    if (mfcc) CU(cudaMemcpyAsync(mfcc + c0 * fpc * 32, s.mfcc, nf * 64, cudaMemcpyDeviceToHost, s.stream));
    if (net_in) CU(cudaMemcpyAsync(net_in + c0 * fpc * EDISON_N_COEFF, s.net_in, nf * EDISON_N_COEFF, cudaMemcpyDeviceToHost, s.stream));
    if (logits) CU(cudaMemcpyAsync(logits + c0 * EDISON_N_CLASSES, s.logits, nc * EDISON_N_CLASSES, cudaMemcpyDeviceToHost, s.stream));
    if (softmax) CU(cudaMemcpyAsync(softmax + c0 * EDISON_N_CLASSES, s.softmax, nc * EDISON_N_CLASSES, cudaMemcpyDeviceToHost, s.stream));
  }
  return EDISON_OK;
}

static int kws_host(const int16_t *pcm, size_t n_clips, size_t fpc, size_t stride, int16_t *mfcc, int8_t *net_in,
                    int8_t *logits, int8_t *softmax) {
  if ((logits || softmax) && fpc != EDISON_N_FRAMES) return fail(EDISON_E_ARG, "inference needs 31 frames per clip");
  return sync_slots(kws_host_chunks(pcm, n_clips, fpc, stride, mfcc, net_in, logits, softmax));
}

static int infer_host_chunks(const int8_t *net_in, size_t n, int8_t *logits, int8_t *softmax) {
  const size_t chunk = (size_t)1 << 18;
  size_t k = 0;
  for (size_t c0 = 0; c0 < n; c0 += chunk, k++) {
    const size_t nc = n - c0 < chunk ? n - c0 : chunk;
    Slot &s = g.slot[k & 1];
    int rc;
    if ((rc = ensure(&s.net_in, &s.net_in_cap, nc * EDISON_NET_IN))) return rc;
    if (logits && (rc = ensure(&s.logits, &s.logits_cap, nc * EDISON_N_CLASSES))) return rc;
    if (softmax && (rc = ensure(&s.softmax, &s.softmax_cap, nc * EDISON_N_CLASSES))) return rc;
    CU(cudaMemcpyAsync(s.net_in, net_in + c0 * EDISON_NET_IN, nc * EDISON_NET_IN, cudaMemcpyHostToDevice, s.stream));
    if ((rc = infer_dev((const int8_t *)s.net_in, nc, logits ? (int8_t *)s.logits : nullptr,
                        softmax ? (int8_t *)s.softmax : nullptr, nullptr, s.stream)))
      return rc;
    if (logits) CU(cudaMemcpyAsync(logits + c0 * EDISON_N_CLASSES, s.logits, nc * EDISON_N_CLASSES, cudaMemcpyDeviceToHost, s.stream));
    if (softmax) CU(cudaMemcpyAsync(softmax + c0 * EDISON_N_CLASSES, s.softmax, nc * EDISON_N_CLASSES, cudaMemcpyDeviceToHost, s.stream));
  }
  return EDISON_OK;
}

int ctx_fail(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof t_err, fmt, ap);
  va_end(ap);
  return code;
}
void ctx_lock(void) { g_mu.lock(); }
void ctx_unlock(void) { g_mu.unlock(); }
int ctx_acquire(CtxView *out) {  // caller holds ctx_lock()
  if (!g.ready) {
    int r = do_init(nullptr);
    if (r) return r;
  }
  out->tables = g.d_tables;
  out->model = g.d_model;
  out->sm_count = g.sm_count;
  out->softmax_mode = g.softmax_mode;
  out->stream = g.slot[0].stream;
  out->data_dir = g.data_dir.c_str();
  out->device = g.device;
  return EDISON_OK;
}

}  // namespace edison

using namespace edison;

#define REQUIRE_READY()                                                                       \
  do {                                                                                        \
    if (!g.ready) {                                                                           \
      int r_ = do_init(nullptr);                                                              \
      if (r_) return r_;                                                                      \
    }                                                                                         \
  } while (0)

extern "C" {

int edison_init(const edison_cfg *cfg) {
  std::lock_guard<std::mutex> lk(g_mu);
  return do_init(cfg);
}

void edison_shutdown(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g.ready) return;
  DeviceScope device_scope_;
  cudaDeviceSynchronize();
  release_device_state();
  mfcc_float_release();
  cube_release();
  g.ready = false;
}

const char *edison_last_error(void) { return t_err; }

int edison_set_softmax_mode(int mode) {
  if (mode < 0 || mode > 2) return fail(EDISON_E_ARG, "softmax_mode %d", mode);
  std::lock_guard<std::mutex> lk(g_mu);
  g.softmax_mode = mode;
  return EDISON_OK;
}

unsigned long long edison_launch_count(void) { return g_launches.load(); }

void *edison_host_alloc(size_t bytes) {
  DeviceScope device_scope_;
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}
void edison_host_free(void *p) {
  if (p) cudaFreeHost(p);
}

int edison_mfcc_q15_clips(const int16_t *pcm, size_t n_clips, size_t frames_per_clip, size_t clip_stride_samples,
                          int16_t *mfcc, int8_t *net_in) {
  std::lock_guard<std::mutex> lk(g_mu);
  REQUIRE_READY();
  DeviceScope device_scope_;
  if (!n_clips || !frames_per_clip) return EDISON_OK;
  if (!pcm || (!mfcc && !net_in)) return fail(EDISON_E_ARG, "null pointer");
  if (n_clips > 1 && clip_stride_samples == 0) return fail(EDISON_E_ARG, "zero clip stride");
  if (int r = same_side(pcm, mfcc, net_in)) return r;
  if (is_device_ptr(pcm)) {
    int r = mfcc_dev(pcm, n_clips, frames_per_clip, clip_stride_samples, mfcc, net_in, g.slot[0].stream);
    if (r) return r;
    CU(cudaStreamSynchronize(g.slot[0].stream));
    return EDISON_OK;
  }
  return kws_host(pcm, n_clips, frames_per_clip, clip_stride_samples, mfcc, net_in, nullptr, nullptr);
}

int edison_mfcc_q15_batch(const int16_t *pcm, size_t n_frames, int16_t *mfcc) {
  return edison_mfcc_q15_clips(pcm, n_frames, 1, EDISON_FRAME_LEN, mfcc, nullptr);
}

int edison_kws_batch(const int16_t *pcm, size_t n_clips, size_t clip_stride_samples, int8_t *logits, int8_t *softmax,
                     int8_t *net_in) {
  std::lock_guard<std::mutex> lk(g_mu);
  REQUIRE_READY();
  DeviceScope device_scope_;
  if (!n_clips) return EDISON_OK;
  if (!pcm || (!logits && !softmax && !net_in)) return fail(EDISON_E_ARG, "null pointer");
  if (n_clips > 1 && clip_stride_samples == 0) return fail(EDISON_E_ARG, "zero clip stride");
  if (int r = same_side(pcm, logits, softmax, net_in)) return r;
  if (is_device_ptr(pcm)) {
    cudaStream_t st = g.slot[0].stream;
    Slot &s = g.slot[0];
    int8_t *d_net = net_in;
    if (!d_net) {
      int r = ensure(&s.net_in, &s.net_in_cap, n_clips * EDISON_NET_IN);
      if (r) return r;
      d_net = (int8_t *)s.net_in;
    }
    int r = mfcc_dev(pcm, n_clips, EDISON_N_FRAMES, clip_stride_samples, nullptr, d_net, st);
    if (r) return r;
    if (logits || softmax) {
      r = infer_dev(d_net, n_clips, logits, softmax, nullptr, st);
      if (r) return r;
    }
    CU(cudaStreamSynchronize(st));
    return EDISON_OK;
  }
  return kws_host(pcm, n_clips, EDISON_N_FRAMES, clip_stride_samples, nullptr, net_in, logits, softmax);
}

int edison_infer_q7_batch(const int8_t *net_in, size_t n, int8_t *logits, int8_t *softmax) {
  std::lock_guard<std::mutex> lk(g_mu);
  REQUIRE_READY();
  DeviceScope device_scope_;
  if (!n) return EDISON_OK;
  if (!net_in || (!logits && !softmax)) return fail(EDISON_E_ARG, "null pointer");
  if (int r = same_side(net_in, logits, softmax)) return r;
  if (is_device_ptr(net_in)) {
    int r = infer_dev(net_in, n, logits, softmax, nullptr, g.slot[0].stream);
    if (r) return r;
    CU(cudaStreamSynchronize(g.slot[0].stream));
    return EDISON_OK;
  }
  return sync_slots(infer_host_chunks(net_in, n, logits, softmax));
}

int edison_mfcc_q15_clips_dev(const int16_t *d_pcm, size_t n_clips, size_t frames_per_clip, size_t clip_stride_samples,
                              int16_t *d_mfcc, int8_t *d_net_in, void *stream) {
  if (!g.ready) return fail(EDISON_E_NOINIT, "edison_init() first");
  DeviceScope device_scope_;
  return mfcc_dev(d_pcm, n_clips, frames_per_clip, clip_stride_samples, d_mfcc, d_net_in,
                  stream ? (cudaStream_t)stream : g.slot[0].stream);
}

int edison_infer_q7_batch_dev(const int8_t *d_net_in, size_t n, int8_t *d_logits, int8_t *d_softmax, void *stream) {
  if (!g.ready) return fail(EDISON_E_NOINIT, "edison_init() first");
  DeviceScope device_scope_;
  return infer_dev(d_net_in, n, d_logits, d_softmax, nullptr, stream ? (cudaStream_t)stream : g.slot[0].stream);
}

int edison_kws_batch_dev(const int16_t *d_pcm, size_t n_clips, size_t clip_stride_samples, int8_t *d_logits,
                         int8_t *d_softmax, int8_t *d_net_in, void *stream) {
  return edison_kws_counts_dev(d_pcm, n_clips, clip_stride_samples, d_logits, d_softmax, d_net_in, nullptr, stream);
}

int edison_kws_counts_dev(const int16_t *d_pcm, size_t n_clips, size_t clip_stride_samples, int8_t *d_logits, int8_t *d_softmax,
                          int8_t *d_net_in, int32_t *d_class_counts, void *stream) {
  if (!g.ready) return fail(EDISON_E_NOINIT, "edison_init() first");
  if (!d_net_in) return fail(EDISON_E_ARG, "d_net_in scratch [n_clips][403] is required by the _dev variant");
  DeviceScope device_scope_;
  cudaStream_t st = stream ? (cudaStream_t)stream : g.slot[0].stream;
  int r = mfcc_dev(d_pcm, n_clips, EDISON_N_FRAMES, clip_stride_samples, nullptr, d_net_in, st);
  if (r) return r;
  if (d_logits || d_softmax || d_class_counts) return infer_dev(d_net_in, n_clips, d_logits, d_softmax, d_class_counts, st);
  return EDISON_OK;
}

int edison_infer_q7_counts_dev(const int8_t *d_net_in, size_t n, int8_t *d_logits, int8_t *d_softmax, int32_t *d_class_counts,
                               void *stream) {
  if (!g.ready) return fail(EDISON_E_NOINIT, "edison_init() first");
  DeviceScope device_scope_;
  return infer_dev(d_net_in, n, d_logits, d_softmax, d_class_counts, stream ? (cudaStream_t)stream : g.slot[0].stream);
}

int edison_sync(void *stream) {
  if (!g.ready) return fail(EDISON_E_NOINIT, "edison_init() first");
  DeviceScope device_scope_;
  CU(cudaStreamSynchronize(stream ? (cudaStream_t)stream : g.slot[0].stream));
  return EDISON_OK;
}

int edison_debug_sqrt_q31(const int32_t *in, size_t n, int32_t *out) {
  std::lock_guard<std::mutex> lk(g_mu);
  REQUIRE_READY();
  DeviceScope device_scope_;
  if (!n) return EDISON_OK;
  int32_t *d_in = nullptr, *d_out = nullptr;
  cudaError_t e = cudaMalloc(&d_in, n * 4);
  if (e == cudaSuccess) e = cudaMalloc(&d_out, n * 4);
  if (e == cudaSuccess) e = cudaMemcpy(d_in, in, n * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = launch_debug_sqrt(d_in, n, d_out, g.slot[0].stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(g.slot[0].stream);
  if (e == cudaSuccess) e = cudaMemcpy(out, d_out, n * 4, cudaMemcpyDeviceToHost);
  if (d_in) cudaFree(d_in);
  if (d_out) cudaFree(d_out);
  if (e != cudaSuccess) return fail(EDISON_E_CUDA, "edison_debug_sqrt_q31: %s", cudaGetErrorString(e));
  return EDISON_OK;
}

int edison_debug_mag_sweep(uint32_t lo, uint32_t hi, unsigned long long *mismatches, unsigned long long *checksum) {
  std::lock_guard<std::mutex> lk(g_mu);
  REQUIRE_READY();
  DeviceScope device_scope_;
  unsigned long long *d = nullptr, h[2] = {0, 0};
  cudaError_t e = cudaMalloc(&d, 16);
  if (e == cudaSuccess) e = cudaMemset(d, 0, 16);
  if (e == cudaSuccess) e = launch_debug_mag_sweep(lo, hi, g.d_tables->sqrt_exc, d, g.slot[0].stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(g.slot[0].stream);
  if (e == cudaSuccess) e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  if (d) cudaFree(d);
  if (e != cudaSuccess) return fail(EDISON_E_CUDA, "edison_debug_mag_sweep: %s", cudaGetErrorString(e));
  if (mismatches) *mismatches = h[0];
  if (checksum) *checksum = h[1];
  return EDISON_OK;
}

// K2 with per-unit clock64() stamps of block 0's second batch: d_trace[4 * units] (device, int64), see nnom_q7_umma.cu
int edison_debug_k2_trace_dev(const int8_t *d_net_in, size_t n, int8_t *d_logits, int8_t *d_softmax, long long *d_trace, void *stream) {
  if (!g.ready) return fail(EDISON_E_NOINIT, "edison_init() first");
  if (!d_trace) return fail(EDISON_E_ARG, "null trace buffer");
  DeviceScope device_scope_;
  K2Launch k{};
  k.net_in = d_net_in; k.n = (long long)n; k.logits = d_logits; k.softmax = d_softmax; k.softmax_mode = g.softmax_mode; k.trace = d_trace;
  CU(launch_nnom_q7(g.d_model, k, g.sm_count, stream ? (cudaStream_t)stream : g.slot[0].stream));
  return EDISON_OK;
}

// used by firmware_api.c
int edison_internal_keyword(uint32_t idx, const char **out) {
  std::lock_guard<std::mutex> lk(g_mu);
  REQUIRE_READY();
  if ((int)idx >= g.n_keywords) {
    *out = nullptr;
    return EDISON_E_ARG;
  }
  *out = g.keywords[idx].c_str();
  return EDISON_OK;
}
int edison_internal_keyword_count(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g.ready && do_init(nullptr)) return 0;
  return g.n_keywords;
}
int edison_internal_input_quant(int *scale, int *mn, int *mx) {
  std::lock_guard<std::mutex> lk(g_mu);
  REQUIRE_READY();
  *scale = g.in_scale;
  *mn = g.in_min;
  *mx = g.in_max;
  return EDISON_OK;
}

}  // extern "C"
