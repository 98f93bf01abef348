// stream.cu -- continuous streaming scheduler: include/edison_b200.h section 3b.
//
// Replaces the firmware's sliding-window loop (appAudioEvent app.c:635-661 + appMicMfccInfereContinuous
// app.c:288-394) for n concurrent streams.  State per stream lives in device memory: the 31x13 int8 window
// (netInput, app.c:71-93) and the 10 float filter values (netOutFilt, app.c:93).  One hop = four launches
// on one stream: K1 over the n new frames -> push kernel (window shift + new row) -> K2 over the n windows
// -> post kernel (moving average, argmax, threshold).
#include <stdio.h>

#include "edison_internal.h"
#include "../../include/edison_b200.h"

struct edison_streams {
  size_t n;
  unsigned long long hops;
  int8_t *d_window;    // [n][403]
  int8_t *d_rows;      // [n][13]  MFCC rows of the current hop
  int8_t *d_softmax;   // [n][10]
  float *d_filt;       // [n][10]
  int32_t *d_detect;   // [n]
  int16_t *d_pcm;      // [n][1024] staging for host callers
  int8_t *d_incr;      // ring state of the incremental net (nnom_q7_kernel.cu), 7,392 B per stream
  edison::NnomIncrState incr;
  int incremental;     // 1: one new row per layer and hop (default); 0: the whole window through K2 every hop
};

namespace edison {

// mfccToNetInputPush (app.c:706-719): rows 1..30 move to 0..29, the new row goes to 30
__global__ void stream_push_kernel(int8_t *window, const int8_t *rows, size_t n) {
  const size_t s = (size_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  if (s >= n) return;
  const int lane = threadIdx.x & 31;
  int8_t *w = window + s * EDISON_NET_IN;
  int8_t v[13];
#pragma unroll
  for (int k = 0; k < 13; k++) {
    const int i = lane + 32 * k;
    v[k] = i < EDISON_NET_IN - EDISON_N_COEFF ? w[i + EDISON_N_COEFF] : (int8_t)0;
  }
  __syncwarp();
#pragma unroll
  for (int k = 0; k < 13; k++) {
    const int i = lane + 32 * k;
    if (i < EDISON_NET_IN - EDISON_N_COEFF) w[i] = v[k];
  }
  if (lane < EDISON_N_COEFF) w[EDISON_NET_IN - EDISON_N_COEFF + lane] = rows[s * EDISON_N_COEFF + lane];
}

// app.c:336-350: netOutFloat = (float)netOutput; netOutFilt = 0.9 netOutFilt + (1.0 - 0.9) netOutFloat, evaluated in
// double (the literals are doubles, the firmware has no -fsingle-precision-constant) and stored as float;
// arm_max_f32 (first maximum); spotted when > TRUE_THRESHOLD 0.5
__global__ void stream_post_kernel(const int8_t *softmax, float *filt, int32_t *detect, size_t n) {
  const size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const double alpha = 0.9, beta = 1.0 - 0.9;
  float best = 0.f;
  int idx = 0;
  for (int i = 0; i < EDISON_N_CLASSES; i++) {
    const double o = (double)(float)softmax[s * EDISON_N_CLASSES + i];
    const float f = (float)__dadd_rn(__dmul_rn(alpha, (double)filt[s * EDISON_N_CLASSES + i]), __dmul_rn(beta, o));
    filt[s * EDISON_N_CLASSES + i] = f;
    if (i == 0 || f > best) { best = f; idx = i; }
  }
  if (detect) detect[s] = best > 0.5f ? idx : -1;
}

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

static int push_dev(edison_streams *s, const CtxView &c, const int16_t *d_pcm, int8_t *d_softmax, float *d_filt,
                    int32_t *d_detect, cudaStream_t st) {
  const size_t n = s->n;
  MfccLaunch L;
  L.pcm = d_pcm; L.n_clips = (long long)n; L.frames_per_clip = 1; L.clip_stride = EDISON_FRAME_LEN;
  L.mfcc = nullptr; L.net_in = s->d_rows; L.tables = c.tables;
  CUS(launch_mfcc_q15(L, c.sm_count, st));
  stream_push_kernel<<<(unsigned)((n + 7) / 8), 256, 0, st>>>(s->d_window, s->d_rows, n);
  count_launch();
  s->hops++;
  // the firmware fills the window (31 frames), waits for one more frame, then starts inferring (app.c:311-318)
  const bool infer = s->hops > EDISON_N_FRAMES;
  if (s->incremental)   // runs from the first hop on: it is what fills the rings
    CUS(launch_nnom_q7_incremental(c.model, s->d_window, s->incr, (long long)n, (long long)s->hops - 1, nullptr,
                                   infer ? s->d_softmax : nullptr, c.softmax_mode, c.sm_count, st));
  if (infer) {
    if (!s->incremental)
      CUS(launch_nnom_q7(c.model, s->d_window, (long long)n, nullptr, s->d_softmax, c.softmax_mode, c.sm_count, st));
    stream_post_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(s->d_softmax, s->d_filt, s->d_detect, n);
    count_launch();
  } else {
    stream_fill_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(s->d_detect, n, -2);
    count_launch();
  }
  CUS(cudaGetLastError());
  if (d_softmax) CUS(cudaMemcpyAsync(d_softmax, s->d_softmax, n * EDISON_N_CLASSES, cudaMemcpyDefault, st));
  if (d_filt) CUS(cudaMemcpyAsync(d_filt, s->d_filt, n * EDISON_N_CLASSES * sizeof(float), cudaMemcpyDefault, st));
  if (d_detect) CUS(cudaMemcpyAsync(d_detect, s->d_detect, n * sizeof(int32_t), cudaMemcpyDefault, st));
  return EDISON_OK;
}

}  // namespace edison

using namespace edison;

extern "C" {

int edison_streams_open(size_t n_streams, edison_streams **out) {
  DeviceScope device_scope_;
  if (!out || !n_streams) return ctx_fail(EDISON_E_ARG, "edison_streams_open: bad arguments");
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  ctx_unlock();
  if (rc) return rc;
  edison_streams *s = new edison_streams();
  s->n = n_streams;
  s->hops = 0;
  cudaError_t e = cudaSuccess;
  if (e == cudaSuccess) e = cudaMalloc(&s->d_window, n_streams * EDISON_NET_IN);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_rows, n_streams * EDISON_N_COEFF);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_softmax, n_streams * EDISON_N_CLASSES);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_filt, n_streams * EDISON_N_CLASSES * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&s->d_detect, n_streams * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&s->d_pcm, n_streams * EDISON_FRAME_LEN * sizeof(int16_t));
  s->incremental = 1;
  if (e == cudaSuccess) e = cudaMalloc(&s->d_incr, nnom_incr_state_bytes((long long)n_streams));
  if (e == cudaSuccess) e = cudaMemset(s->d_incr, 0, nnom_incr_state_bytes((long long)n_streams));
  if (e == cudaSuccess) nnom_incr_state_bind(&s->incr, s->d_incr, (long long)n_streams);
  if (e == cudaSuccess) e = cudaMemset(s->d_window, 0, n_streams * EDISON_NET_IN);   // netInput is zero-initialised static storage
  if (e == cudaSuccess) e = cudaMemset(s->d_filt, 0, n_streams * EDISON_N_CLASSES * sizeof(float));  // app.c:300
  if (e == cudaSuccess) e = cudaMemset(s->d_softmax, 0, n_streams * EDISON_N_CLASSES);
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
  void *bufs[] = {s->d_window, s->d_rows, s->d_softmax, s->d_filt, s->d_detect, s->d_pcm, s->d_incr};
  for (void *b : bufs)
    if (b) cudaFree(b);
  delete s;
}

unsigned long long edison_streams_hops(const edison_streams *s) { return s ? s->hops : 0; }

int edison_streams_set_incremental(edison_streams *s, int on) {
  if (!s) return ctx_fail(EDISON_E_ARG, "edison_streams_set_incremental: null handle");
  if (s->hops) return ctx_fail(EDISON_E_ARG, "edison_streams_set_incremental: only before the first hop (the rings fill from hop 1)");
  s->incremental = on ? 1 : 0;
  return EDISON_OK;
}

int edison_streams_push_hop_dev(edison_streams *s, const int16_t *d_hop_pcm, int8_t *d_softmax, float *d_filt,
                                int32_t *d_detect, void *stream) {
  DeviceScope device_scope_;
  if (!s || !d_hop_pcm) return ctx_fail(EDISON_E_ARG, "edison_streams_push_hop_dev: null pointer");
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  if (!rc) rc = push_dev(s, c, d_hop_pcm, d_softmax, d_filt, d_detect, stream ? (cudaStream_t)stream : c.stream);
  ctx_unlock();
  return rc;
}

int edison_streams_push_hop(edison_streams *s, const int16_t *hop_pcm, int8_t *softmax, float *filt, int32_t *detect) {
  DeviceScope device_scope_;
  if (!s || !hop_pcm) return ctx_fail(EDISON_E_ARG, "edison_streams_push_hop: null pointer");
  ctx_lock();
  CtxView c;
  int rc = ctx_acquire(&c);
  if (!rc) {
    const int16_t *d_pcm = hop_pcm;
    cudaError_t e = cudaSuccess;
    if (!dev_ptr(hop_pcm)) {
      e = cudaMemcpyAsync(s->d_pcm, hop_pcm, s->n * EDISON_FRAME_LEN * sizeof(int16_t), cudaMemcpyHostToDevice, c.stream);
      d_pcm = s->d_pcm;
    }
    if (e != cudaSuccess) rc = ctx_fail(EDISON_E_CUDA, "hop upload: %s", cudaGetErrorString(e));
    if (!rc) rc = push_dev(s, c, d_pcm, softmax, filt, detect, c.stream);
    if (!rc) {
      e = cudaStreamSynchronize(c.stream);
      if (e != cudaSuccess) rc = ctx_fail(EDISON_E_CUDA, "hop sync: %s", cudaGetErrorString(e));
    }
  }
  ctx_unlock();
  return rc;
}

int edison_streams_window(edison_streams *s, size_t stream_index, int8_t *window403) {
  DeviceScope device_scope_;
  if (!s || !window403 || stream_index >= s->n) return ctx_fail(EDISON_E_ARG, "edison_streams_window: bad arguments");
  cudaError_t e = cudaDeviceSynchronize();
  if (e == cudaSuccess) e = cudaMemcpy(window403, s->d_window + stream_index * EDISON_NET_IN, EDISON_NET_IN, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "edison_streams_window: %s", cudaGetErrorString(e));
  return EDISON_OK;
}

}  // extern "C"
