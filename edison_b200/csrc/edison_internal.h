// edison_internal.h -- declarations shared by the translation units of libedison_b200.so
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "mfcc_q15_core.cuh"

namespace edison {

// ---- front-end ----------------------------------------------------------------------------------
struct MfccLaunch {
  const int16_t *pcm;          // device
  long long n_clips;
  int frames_per_clip;
  long long clip_stride;       // samples
  int16_t *mfcc;               // device [n_clips*frames_per_clip][32] or nullptr
  int8_t *net_in;              // device: 13 int8 per frame at net_in + frame * net_in_stride, or nullptr
  long long net_in_stride;     // bytes between consecutive frames' rows: 13 = packed [n_frames][13]; 403 = one row of n windows
  const MfccTables *tables;    // device
  int group_frames = 8;        // set by launch_mfcc_q15: frames per warp group (8; 1, 2 or 4 for small launches)
};
// fills the twiddle / split / sqrt-exception tables (formula-generated, checksummed against the
// values of the reference's CMSIS build) and copies the mel blob in; returns 0 or EDISON_E_*
int build_mfcc_tables(MfccTables *host_tables, const char *mel_blob_path, char *err, size_t errlen);
cudaError_t launch_mfcc_q15(const MfccLaunch &p, int sm_count, cudaStream_t stream);
cudaError_t mfcc_q15_configure(void);   // opt-in to large dynamic shared memory (once)

// ---- int8 net -----------------------------------------------------------------------------------
struct NnomModel;   // opaque handle: the device pointer of K2's weight image (nnom_q7_kernel.cu)
int nnom_load(const char *blob_path, NnomModel **out_dev_model, int *in_scale, int *in_min, int *in_max,
              char *err, size_t errlen);
void nnom_free(NnomModel *m);
cudaError_t nnom_configure(void);
// K2 (tcgen05 kernel, nnom_q7_umma.cu)
struct K2Launch {
  const int8_t *net_in;     // device [n][403]; row r of window i is at net_in + i * 403 + ((r + row0) % 31) * 13
  long long n;
  int8_t *logits, *softmax; // device [n][10], either may be null
  int softmax_mode;
  int *class_counts;        // optional device int32[10], += number of clips per first-maximum logit class
  int row0;                 // 0 for plain batches; the streaming windows are rings of rows (stream.cu)
  float *filt;              // optional streaming post-processing fused into the dense epilogue (app.c:336-350):
  int32_t *detect;          //   filt[n][10] moving average (in / out), detect[n] = spotted class or -1
  long long *trace;         // debug: clock64() stamps of block 0 (edison_debug_k2_trace_dev)
  // incremental sliding-window mode (k2_umma_plan.h): the row cache [batches][CACHE_BATCH_BYTES], the phase set of this hop and the
  // slots holding the rows of the set's two previous uses / taking this use's rows; non-null selects the incremental kernel
  uint8_t *cache;
  int cache_set, cache_older, cache_newer, cache_write;
};
size_t nnom_umma_cache_bytes(long long n);   // bytes of the row cache for n streams
cudaError_t launch_nnom_q7(const NnomModel *model, const K2Launch &k, int sm_count, cudaStream_t stream);
cudaError_t nnom_umma_configure(void);
int nnom_umma_upload(const int8_t *const w[5], const int8_t *const b[5], const int rs[5], const int ls[5], uint8_t **out_image);
cudaError_t launch_nnom_q7_umma(const uint8_t *image, const K2Launch &k, int sm_count, cudaStream_t stream);

// ---- debug kernels --------------------------------------------------------------------------------
cudaError_t launch_debug_sqrt(const int32_t *d_in, size_t n, int32_t *d_out, cudaStream_t stream);
cudaError_t launch_debug_mag_sweep(uint32_t lo, uint32_t hi, const uint32_t *d_exc, unsigned long long *d_out2,
                                   cudaStream_t stream);

void count_launch(void);

// NVTX ranges named after the reference's cycle-profiler labels (prfStart / prfEvent: firmware/src/audioprocessing.c:123-207,
// firmware/src/app.c:177-207), so that a timeline tool shows the same stage names a firmware trace does.  Header-only NVTX v3:
// a no-op unless a profiler injects itself.  The ranges cover the host-side enqueue of the corresponding work (the stages inside
// audioCalcMFCCs -- "fft", "cmpl_mag_q15", "matrix mult", "dct2_q15" -- are one fused kernel here).
struct TraceRange {
  explicit TraceRange(const char *label);
  ~TraceRange();
  TraceRange(const TraceRange &) = delete;
  TraceRange &operator=(const TraceRange &) = delete;
};
void mfcc_float_release(void);   // frees the lazily built tables of K3 (mfcc_float.cu); caller holds the ctx lock
void cube_release(void);         // frees the float model of K4 (cube_f32_kernel.cu)

// ---- shared context (edison_api.cu) for the other translation units of the C ABI ----------------------
struct CtxView {
  const MfccTables *tables;   // device
  const NnomModel *model;     // device
  int sm_count;
  int softmax_mode;
  cudaStream_t stream;        // the library's default stream
  const char *data_dir;       // directory of the model / table blobs (valid until edison_shutdown)
  int device;                 // the CUDA device the library is bound to
};
// One context per process, bound to one device (edison_init).  Every public entry point makes that device current for the
// calling thread and restores the caller's on return, so calls from other threads or after the caller switched devices
// still allocate, copy and launch on the library's GPU.
int ctx_device(void);                              // the bound device, -1 before edison_init
struct DeviceScope {
  int prev = -1, want = -1;
  DeviceScope() {
    want = ctx_device();
    if (want >= 0 && cudaGetDevice(&prev) == cudaSuccess && prev != want) cudaSetDevice(want);
    else prev = want;
  }
  ~DeviceScope() {
    if (prev != want && prev >= 0) cudaSetDevice(prev);
  }
  DeviceScope(const DeviceScope &) = delete;
  DeviceScope &operator=(const DeviceScope &) = delete;
};
int ctx_acquire(CtxView *out);                     // initialises the library on first use; EDISON_OK or an error code
int ctx_fail(int code, const char *fmt, ...);      // records the thread's last-error text, returns code
void ctx_lock(void);
void ctx_unlock(void);

}  // namespace edison
