// stream.cu -- continuous streaming scheduler: include/edison_b200.h section 3b.
//
// Replaces the firmware's sliding-window loop (appAudioEvent app.c:635-661 + appMicMfccInfereContinuous app.c:288-394) for n
// concurrent streams.  State per stream lives in device memory: the 31x13 int8 window (netInput, app.c:71-93) and the 10 float
// filter values (netOutFilt, app.c:93).
//
// One hop = TWO launches on one stream:
//   K1 over the n new frames, storing each stream's 13-coefficient row straight into its window.  The window is a RING of rows
//      (logical row r at physical row (r + row0) % 31), so mfccToNetInputPush (app.c:706-719: shift 30 rows up, append) is the
//      store of one row plus row0 <- row0 + 1: no shift kernel, no 2 x 40 MB of window traffic per 100,000-stream hop;
//   K2 over the n windows (tcgen05 kernel; its input layout step applies the ring offset) with the firmware's post-processing
//      (moving average in double, first-maximum argmax, 0.5 threshold; app.c:336-350) fused into its dense epilogue.
// (Round 1 issued four launches per hop -- K1, push, K2, post -- and a 1,024-stream hop was launch-latency bound.)
// While the window fills (hops 1..31, like the firmware: app.c:311-318) a hop is K1 plus a tiny fill of detect[] = -2.
// Incremental mode (edison_streams_set_incremental): K2's incremental kernel instead -- only the bottom row of every layer is computed per hop, the
// older rows come from a row cache in HBM (k2_umma_plan.h); it runs from the first hop on (that is what fills the cache), same two launches, same bytes.
//
// Host callers: edison_streams_submit_hop / edison_streams_collect_hop pipeline hops through two staging slots -- the H2D copy of
// hop h + 1 runs on a copy stream while hop h computes, results come back through pinned staging -- and edison_streams_push_hop is
// submit + collect.
#include <stdio.h>
#include <string.h>

#include "edison_internal.h"
#include "../../include/edison_b200.h"

struct edison_streams {
  size_t n;
  unsigned long long hops;        // hops submitted
  unsigned long long collected;   // hops whose results the host caller has taken (submit / collect API)
  int row0;                       // physical row of the window's logical row 0 (the oldest row)
  int8_t *d_window;    // [n][31 physical rows][13]
  int8_t *d_softmax;   // [n][10]
  float *d_filt;       // [n][10]
  int32_t *d_detect;   // [n]
  int incremental;     // 1 (default): only the bottom row of every layer per hop, older rows from the row cache (nnom_q7_umma_incr_kernel;
                       // bit-identical, faster, 6.5 KB of HBM per stream); 0: the whole window through K2 every hop, as the firmware does
  uint8_t *d_cache;    // incremental mode: row cache of nnom_q7_umma.cu (k2_umma_plan.h)
  long long *d_trace;  // debug: K2's per-unit clock stamps of the next hops (edison_streams_debug_trace)
  // host ingest pipeline
  int16_t *d_pcm[2];         // [n][1024] staging
  unsigned char *h_out[2];   // pinned: softmax [n][10] | filt [n][10] float | detect [n] int32
  cudaStream_t compute, copy;
  cudaEvent_t ev_copied[2], ev_computed[2], ev_done[2];
};

namespace edison {

__global__ void stream_fill_kernel(int32_t *detect, size_t n, int32_t v) {
  const size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n) detect[s] = v;
}

static bool dev_ptr(const void *p) {
  cudaPointerAttributes a;
  if (!p || cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

#define CUS(call)                                                                                      \
  do {                                                                                                 \
    cudaError_t e_ = (call);                                                                           \
    if (e_ != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "%s: %s", #call, cudaGetErrorString(e_));    \
  } while (0)

// one hop from device PCM; outputs (device pointers, any may be null) are copies of the streams' state after the hop
static int push_dev(edison_streams *s, const CtxView &c, const int16_t *d_pcm, int8_t *d_softmax, float *d_filt, int32_t *d_detect,
                    cudaStream_t st) {
  const size_t n = s->n;
  // mfccToNetInputPush: the oldest row's slot takes the new row, then the ring advances
  MfccLaunch L;
  L.pcm = d_pcm; L.n_clips = (long long)n; L.frames_per_clip = 1; L.clip_stride = EDISON_FRAME_LEN;
  L.mfcc = nullptr; L.net_in = s->d_window + s->row0 * EDISON_N_COEFF; L.net_in_stride = EDISON_NET_IN; L.tables = c.tables;
  CUS(launch_mfcc_q15(L, c.sm_count, st));
  s->row0 = (s->row0 + 1) % EDISON_N_FRAMES;
  s->hops++;
  // the firmware fills the window (31 frames), waits for one more frame, then starts inferring (app.c:311-318)
  const bool infer = s->hops > EDISON_N_FRAMES;
  if (s->incremental) {   // runs from the first hop on: it is what fills the row cache
    const unsigned long long h = s->hops;
    const int use = (int)((h >> 2) % 3);
    K2Launch k{};
    k.net_in = s->d_window; k.n = (long long)n; k.softmax_mode = c.softmax_mode; k.row0 = s->row0;
    k.cache = s->d_cache; k.cache_set = (int)(h & 3); k.cache_write = use; k.cache_newer = (use + 2) % 3; k.cache_older = (use + 1) % 3;
    if (infer) { k.softmax = s->d_softmax; k.filt = s->d_filt; k.detect = s->d_detect; }
    k.trace = s->d_trace;
    CUS(launch_nnom_q7(c.model, k, c.sm_count, st));
  } else if (infer) {
    K2Launch k{};
    k.net_in = s->d_window; k.n = (long long)n; k.softmax = s->d_softmax; k.softmax_mode = c.softmax_mode; k.row0 = s->row0;
    k.filt = s->d_filt; k.detect = s->d_detect; k.trace = s->d_trace;
    CUS(launch_nnom_q7(c.model, k, c.sm_count, st));
  }
  if (!infer) {
    stream_fill_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(s->d_detect, n, -2);
    count_launch();
  }
  CUS(cudaGetLastError());
  if (d_softmax) CUS(cudaMemcpyAsync(d_softmax, s->d_softmax, n * EDISON_N_CLASSES, cudaMemcpyDefault, st));
  if (d_filt) CUS(cudaMemcpyAsync(d_filt, s->d_filt, n * EDISON_N_CLASSES * sizeof(float), cudaMemcpyDefault, st));
  if (d_detect) CUS(cudaMemcpyAsync(d_detect, s->d_detect, n * sizeof(int32_t), cudaMemcpyDefault, st));
  return EDISON_OK;
}

static size_t out_bytes(const edison_streams *s) { return s->n * (EDISON_N_CLASSES + EDISON_N_CLASSES * sizeof(float) + sizeof(int32_t)); }

}  // namespace edison

using namespace edison;

extern "C" {

int edison_streams_open(size_t n_streams, edison_streams **out) {
  if (!out || !n_streams) return ctx_fail(EDISON_E_ARG, "edison_streams_open: bad arguments");
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  ctx_unlock();
  if (rc) return rc;
  DeviceScope device_scope_;   // after ctx_acquire: the first call may just have initialised the library
  edison_streams *s = new edison_streams();
  memset(s, 0, sizeof *s);
  s->n = n_streams;
  cudaError_t e = cudaSuccess;
  if (e == cudaSuccess) e = cudaMalloc(&s->d_window, n_streams * EDISON_NET_IN);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_softmax, n_streams * EDISON_N_CLASSES);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_filt, n_streams * EDISON_N_CLASSES * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&s->d_detect, n_streams * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMemset(s->d_window, 0, n_streams * EDISON_NET_IN);   // netInput is zero-initialised static storage
  if (e == cudaSuccess) e = cudaMemset(s->d_filt, 0, n_streams * EDISON_N_CLASSES * sizeof(float));  // app.c:300
  if (e == cudaSuccess) e = cudaMemset(s->d_softmax, 0, n_streams * EDISON_N_CLASSES);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_cache, nnom_umma_cache_bytes((long long)n_streams));   // incremental mode is the default
  if (e == cudaSuccess) e = cudaMemset(s->d_cache, 0, nnom_umma_cache_bytes((long long)n_streams));
  s->incremental = 1;
  if (e == cudaSuccess) e = cudaDeviceSynchronize();   // the memsets ran on the legacy stream; the hops run on non-blocking streams that do not wait for it
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->compute, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->copy, cudaStreamNonBlocking);
  for (int i = 0; i < 2 && e == cudaSuccess; i++) {
    e = cudaEventCreateWithFlags(&s->ev_copied[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s->ev_computed[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s->ev_done[i], cudaEventDisableTiming);
  }
  if (e != cudaSuccess) {
    edison_streams_close(s);
    return ctx_fail(EDISON_E_NOMEM, "edison_streams_open(%zu): %s", n_streams, cudaGetErrorString(e));
  }
  *out = s;
  return EDISON_OK;
}

void edison_streams_close(edison_streams *s) {
  DeviceScope device_scope_;
  if (!s) return;
  cudaDeviceSynchronize();
  void *bufs[] = {s->d_window, s->d_softmax, s->d_filt, s->d_detect, s->d_pcm[0], s->d_pcm[1], s->d_cache};
  for (void *b : bufs)
    if (b) cudaFree(b);
  for (int i = 0; i < 2; i++) {
    if (s->h_out[i]) cudaFreeHost(s->h_out[i]);
    if (s->ev_copied[i]) cudaEventDestroy(s->ev_copied[i]);
    if (s->ev_computed[i]) cudaEventDestroy(s->ev_computed[i]);
    if (s->ev_done[i]) cudaEventDestroy(s->ev_done[i]);
  }
  if (s->compute) cudaStreamDestroy(s->compute);
  if (s->copy) cudaStreamDestroy(s->copy);
  delete s;
}

unsigned long long edison_streams_hops(const edison_streams *s) { return s ? s->hops : 0; }

// debug: K2 of the following hops runs its tracing instantiation and writes block 0's per-unit clock stamps to d_trace (128 int64, device); null: off
int edison_streams_debug_trace(edison_streams *s, long long *d_trace) {
  if (!s) return ctx_fail(EDISON_E_ARG, "edison_streams_debug_trace: null handle");
  s->d_trace = d_trace;
  return EDISON_OK;
}

int edison_streams_set_incremental(edison_streams *s, int on) {
  DeviceScope device_scope_;
  if (!s) return ctx_fail(EDISON_E_ARG, "edison_streams_set_incremental: null handle");
  if (s->hops) return ctx_fail(EDISON_E_ARG, "edison_streams_set_incremental: only before the first hop (the row cache fills from hop 1)");
  if (on && !s->d_cache) {
    const size_t bytes = nnom_umma_cache_bytes((long long)s->n);
    cudaError_t e = cudaMalloc(&s->d_cache, bytes);
    if (e == cudaSuccess) e = cudaMemset(s->d_cache, 0, bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return ctx_fail(EDISON_E_NOMEM, "edison_streams_set_incremental: %s", cudaGetErrorString(e));
  }
  if (!on && s->d_cache) { cudaFree(s->d_cache); s->d_cache = nullptr; }
  s->incremental = on ? 1 : 0;
  return EDISON_OK;
}

int edison_streams_push_hop_dev(edison_streams *s, const int16_t *d_hop_pcm, int8_t *d_softmax, float *d_filt,
                                int32_t *d_detect, void *stream) {
  DeviceScope device_scope_;
  if (!s || !d_hop_pcm) return ctx_fail(EDISON_E_ARG, "edison_streams_push_hop_dev: null pointer");
  if (s->hops != s->collected) return ctx_fail(EDISON_E_ARG, "edison_streams_push_hop_dev: a submitted hop has not been collected");
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  if (!rc) rc = push_dev(s, c, d_hop_pcm, d_softmax, d_filt, d_detect, stream ? (cudaStream_t)stream : s->compute);
  if (!rc) s->collected = s->hops;
  ctx_unlock();
  return rc;
}

int edison_streams_submit_hop(edison_streams *s, const int16_t *hop_pcm) {
  DeviceScope device_scope_;
  if (!s || !hop_pcm) return ctx_fail(EDISON_E_ARG, "edison_streams_submit_hop: null pointer");
  if (s->hops - s->collected >= 2) return ctx_fail(EDISON_E_ARG, "edison_streams_submit_hop: two hops in flight, collect one first");
  const int slot = (int)(s->hops & 1);
  const size_t pcm_bytes = s->n * EDISON_FRAME_LEN * sizeof(int16_t);
  for (int i = 0; i < 2; i++) {   // staging, on first use
    if (!s->d_pcm[i] && cudaMalloc(&s->d_pcm[i], pcm_bytes) != cudaSuccess) return ctx_fail(EDISON_E_NOMEM, "hop staging");
    if (!s->h_out[i] && cudaMallocHost(&s->h_out[i], out_bytes(s)) != cudaSuccess) return ctx_fail(EDISON_E_NOMEM, "pinned result staging");
  }
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  if (!rc) {
    const int16_t *d_pcm = hop_pcm;
    cudaError_t e = cudaSuccess;
    if (!dev_ptr(hop_pcm)) {
      // slot reuse: the hop that last read this staging buffer (two hops ago) has computed
      if (s->hops >= 2) e = cudaStreamWaitEvent(s->copy, s->ev_computed[slot], 0);
      if (e == cudaSuccess) e = cudaMemcpyAsync(s->d_pcm[slot], hop_pcm, pcm_bytes, cudaMemcpyHostToDevice, s->copy);
      if (e == cudaSuccess) e = cudaEventRecord(s->ev_copied[slot], s->copy);
      if (e == cudaSuccess) e = cudaStreamWaitEvent(s->compute, s->ev_copied[slot], 0);
      d_pcm = s->d_pcm[slot];
    }
    if (e != cudaSuccess) rc = ctx_fail(EDISON_E_CUDA, "hop upload: %s", cudaGetErrorString(e));
    if (!rc) rc = push_dev(s, c, d_pcm, nullptr, nullptr, nullptr, s->compute);
    if (!rc) {
      unsigned char *h = s->h_out[slot];
      const size_t n = s->n;
      e = cudaEventRecord(s->ev_computed[slot], s->compute);
      if (e == cudaSuccess) e = cudaMemcpyAsync(h, s->d_softmax, n * EDISON_N_CLASSES, cudaMemcpyDeviceToHost, s->compute);
      if (e == cudaSuccess)
        e = cudaMemcpyAsync(h + n * EDISON_N_CLASSES, s->d_filt, n * EDISON_N_CLASSES * sizeof(float), cudaMemcpyDeviceToHost, s->compute);
      if (e == cudaSuccess)
        e = cudaMemcpyAsync(h + n * EDISON_N_CLASSES * (1 + sizeof(float)), s->d_detect, n * sizeof(int32_t), cudaMemcpyDeviceToHost, s->compute);
      if (e == cudaSuccess) e = cudaEventRecord(s->ev_done[slot], s->compute);
      if (e != cudaSuccess) rc = ctx_fail(EDISON_E_CUDA, "hop results: %s", cudaGetErrorString(e));
    }
  }
  ctx_unlock();
  return rc;
}

int edison_streams_collect_hop(edison_streams *s, int8_t *softmax, float *filt, int32_t *detect) {
  DeviceScope device_scope_;
  if (!s) return ctx_fail(EDISON_E_ARG, "edison_streams_collect_hop: null handle");
  if (s->collected >= s->hops) return ctx_fail(EDISON_E_ARG, "edison_streams_collect_hop: no hop in flight");
  const int slot = (int)(s->collected & 1);
  cudaError_t e = cudaEventSynchronize(s->ev_done[slot]);
  if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "hop sync: %s", cudaGetErrorString(e));
  const unsigned char *h = s->h_out[slot];
  const size_t n = s->n;
  // the caller's result buffers may be host or device memory
  if (softmax) e = cudaMemcpy(softmax, h, n * EDISON_N_CLASSES, cudaMemcpyDefault);
  if (filt && e == cudaSuccess) e = cudaMemcpy(filt, h + n * EDISON_N_CLASSES, n * EDISON_N_CLASSES * sizeof(float), cudaMemcpyDefault);
  if (detect && e == cudaSuccess) e = cudaMemcpy(detect, h + n * EDISON_N_CLASSES * (1 + sizeof(float)), n * sizeof(int32_t), cudaMemcpyDefault);
  if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "hop results: %s", cudaGetErrorString(e));
  s->collected++;
  return EDISON_OK;
}

int edison_streams_push_hop(edison_streams *s, const int16_t *hop_pcm, int8_t *softmax, float *filt, int32_t *detect) {
  if (!s || !hop_pcm) return ctx_fail(EDISON_E_ARG, "edison_streams_push_hop: null pointer");
  if (s->hops != s->collected) return ctx_fail(EDISON_E_ARG, "edison_streams_push_hop: collect the submitted hops first");
  int rc = edison_streams_submit_hop(s, hop_pcm);
  if (rc) return rc;
  return edison_streams_collect_hop(s, softmax, filt, detect);
}

int edison_streams_window(edison_streams *s, size_t stream_index, int8_t *window403) {
  DeviceScope device_scope_;
  if (!s || !window403 || stream_index >= s->n) return ctx_fail(EDISON_E_ARG, "edison_streams_window: bad arguments");
  int8_t ring[EDISON_NET_IN];
  cudaError_t e = cudaDeviceSynchronize();
  if (e == cudaSuccess) e = cudaMemcpy(ring, s->d_window + stream_index * EDISON_NET_IN, EDISON_NET_IN, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "edison_streams_window: %s", cudaGetErrorString(e));
  for (int r = 0; r < EDISON_N_FRAMES; r++)   // logical row r is physical row (r + row0) % 31
    memcpy(window403 + r * EDISON_N_COEFF, ring + ((r + s->row0) % EDISON_N_FRAMES) * EDISON_N_COEFF, EDISON_N_COEFF);
  return EDISON_OK;
}

}  // extern "C"
