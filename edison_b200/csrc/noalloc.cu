// noalloc.cu -- edison_kws_batch_noalloc_dev: K1 -> K2 over a batch of clips with no allocation on the call path.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  edison_kws_batch_dev makes the caller bring the
// [n_clips][403] int8 network-input scratch that K1 writes and K2 reads; here that scratch is a file-scope device array of
// EDISON_NOALLOC_MAX_CLIPS x 403 bytes (3.15 MiB) that every GPU loading the library holds until the process exits.
//
//   * In host code `g_scratch` is a host-side stand-in for the symbol, not a device address.  The address the kernels get comes
//     from cudaGetSymbolAddress, looked up once per device with that device current (scratch_for_device).
//   * Nothing in the scratch has to start at zero: K1 stores all 403 bytes of each of the n_clips rows before K2 reads them, and
//     K2 never reads a row >= n_clips (a ragged last batch goes through its guarded path, nnom_q7_umma.cu transpose_in).
//   * All calls share the one array: two calls whose kernels may overlap on the GPU (different streams, no dependency) race.
//   * The call path below performs no cudaMalloc / cudaFree / cudaMallocHost and calls no allocating library routine; the model
//     image and the MFCC tables it uses were uploaded by edison_init, which is why it refuses to initialise lazily.
#include <stdint.h>

#include "edison_internal.h"
#include "../../include/edison_b200.h"

namespace edison {
namespace {

static __device__ __align__(256) int8_t g_scratch[(size_t)EDISON_NOALLOC_MAX_CLIPS * EDISON_NET_IN];

constexpr int kMaxDevices = 64;
int8_t *g_scratch_addr[kMaxDevices];   // per device, filled on first use (under the ctx lock)

// the device address of g_scratch on `device`, which must be current
int scratch_for_device(int device, int8_t **out) {
  if (device < 0 || device >= kMaxDevices) return ctx_fail(EDISON_E_ARG, "device %d out of range", device);
  if (!g_scratch_addr[device]) {
    void *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return ctx_fail(EDISON_E_CUDA, "cudaGetSymbolAddress(scratch) on device %d: %s", device, cudaGetErrorString(e));
    }
    g_scratch_addr[device] = static_cast<int8_t *>(p);
  }
  *out = g_scratch_addr[device];
  return EDISON_OK;
}

}  // namespace
}  // namespace edison

using namespace edison;

extern "C" int edison_kws_batch_noalloc_dev(const int16_t *d_pcm, size_t n_clips, size_t clip_stride_samples, int8_t *d_logits,
                                            int8_t *d_softmax, void *stream) {
  if (ctx_device() < 0) return ctx_fail(EDISON_E_NOINIT, "edison_init() first (this entry point never allocates, so it cannot initialise)");
  if (n_clips > (size_t)EDISON_NOALLOC_MAX_CLIPS)
    return ctx_fail(EDISON_E_ARG, "edison_kws_batch_noalloc_dev: %zu clips, the static scratch holds at most %d; split the batch or use edison_kws_batch_dev",
                    n_clips, EDISON_NOALLOC_MAX_CLIPS);
  if (!n_clips) return EDISON_OK;
  if (!d_pcm || (!d_logits && !d_softmax)) return ctx_fail(EDISON_E_ARG, "null pointer");
  if (n_clips > 1 && clip_stride_samples == 0) return ctx_fail(EDISON_E_ARG, "zero clip stride");
  if (reinterpret_cast<uintptr_t>(d_pcm) & 1u) return ctx_fail(EDISON_E_ARG, "pcm must be 2-byte aligned");

  DeviceScope device_scope_;
  CtxView ctx;
  int8_t *scratch = nullptr;
  ctx_lock();
  int r = ctx_acquire(&ctx);   // initialised already (checked above): only reads the context
  if (!r) r = scratch_for_device(ctx.device, &scratch);
  ctx_unlock();
  if (r) return r;
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx.stream;

  MfccLaunch L;
  L.pcm = d_pcm; L.n_clips = (long long)n_clips; L.frames_per_clip = EDISON_N_FRAMES; L.clip_stride = (long long)clip_stride_samples;
  L.mfcc = nullptr; L.net_in = scratch; L.net_in_stride = EDISON_N_COEFF; L.tables = ctx.tables;
  cudaError_t e = launch_mfcc_q15(L, ctx.sm_count, st);
  if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "K1 launch: %s", cudaGetErrorString(e));
  K2Launch k{};
  k.net_in = scratch; k.n = (long long)n_clips; k.logits = d_logits; k.softmax = d_softmax; k.softmax_mode = ctx.softmax_mode;
  e = launch_nnom_q7(ctx.model, k, ctx.sm_count, st);
  if (e != cudaSuccess) return ctx_fail(EDISON_E_CUDA, "K2 launch: %s", cudaGetErrorString(e));
  return EDISON_OK;
}
