/* edison_b200.h -- C ABI of libedison_b200.so
 *
 * A drop-in for the keyword-spotting hot path of huettern/edison: the firmware's
 * own entry points keep their names and signatures (section 1), and batched entry points
 * (section 2) expose the same arithmetic over many frames / clips at once.  Everything behind
 * this header runs as CUDA kernels for sm_100a; there is no CPU implementation in the library
 * and every call fails with EDISON_E_NOGPU when no device is usable.
 *
 * Plain C, plain pointers and sizes: bind it from C, cgo, ctypes, JNI, ....
 * File:line references are relative to the reference tree.
 */
#ifndef EDISON_B200_H
#define EDISON_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- constants of the path ------------------------------------------------------------------ */
#define EDISON_FRAME_LEN      1024  /* AUD_MFCC_FRAME_LEN, firmware/src/audioprocessing.h:14   */
#define EDISON_N_MEL          32    /* MEL_MTX_COLS, firmware/src/audio/mel_constants.h:10      */
#define EDISON_N_COEFF        13    /* aiGetInputShape x, firmware/src/ai/ai.c:205-216          */
#define EDISON_N_FRAMES       31    /* aiGetInputShape y                                        */
#define EDISON_NET_IN         (EDISON_N_FRAMES * EDISON_N_COEFF)   /* 403                       */
#define EDISON_N_CLASSES      10    /* firmware/src/ai/nnom/kws_nnom/keywords.txt               */
#define EDISON_CLIP_SAMPLES   (EDISON_N_FRAMES * EDISON_FRAME_LEN) /* 31,744        */

/* ---- status codes (0 = success, negative = failure; aiRunInference keeps nnom_status_t's
 *      convention 0 == NN_SUCCESS, firmware/src/ai/nnom/inc/nnom.h) ---------------------------- */
#define EDISON_OK             0
#define EDISON_E_ARG         (-1)   /* null pointer, zero stride, misaligned device pointer      */
#define EDISON_E_NOGPU       (-2)   /* no CUDA device / driver: the library has no CPU path      */
#define EDISON_E_CUDA        (-3)   /* a CUDA call failed; see edison_last_error()               */
#define EDISON_E_DATA        (-4)   /* model / table blob missing or malformed                   */
#define EDISON_E_NOINIT      (-5)   /* edison_init() has not succeeded                           */
#define EDISON_E_MODEL       (-6)   /* blob describes a graph this build has no kernels for      */
#define EDISON_E_NOMEM       (-7)

/* softmax variants of the reference builds */
#define EDISON_SOFTMAX_CMSIS_REF   0  /* arm_softmax_q7 reference branch, arm_softmax_q7.c:217-254 */
#define EDISON_SOFTMAX_CMSIS_SIMD  1  /* arm_softmax_q7 SIMD branch :83-215 (firmware build)       */
#define EDISON_SOFTMAX_NNOM_LOCAL  2  /* local_softmax_q7, nnom_local.c:948-1003                   */

typedef struct edison_cfg {
  int device;             /* CUDA device ordinal; -1 = keep the current device                   */
  const char *data_dir;   /* holds mel_q15.bin and kws_nnom_q7.bin; NULL = $EDISON_B200_DATA,
                             else <directory of the .so>/data                                */
  int softmax_mode;       /* EDISON_SOFTMAX_*; edison_init(NULL) selects EDISON_SOFTMAX_CMSIS_SIMD,
                             the firmware build's variant                                        */
  size_t chunk_clips;     /* host-pointer calls are pipelined in chunks of this many clips
                             (0 = default 8192)                                                  */
} edison_cfg;

/* ---- section 0: library lifetime -------------------------------------------------------------- */
/* One context per process, bound to one device.  cfg may be NULL (current device, default data directory, the firmware
 * build's softmax).  A second call returns EDISON_OK when it asks for what is already there and EDISON_E_ARG when it asks
 * for another device, softmax mode or data directory (edison_shutdown() first).  Every entry point makes the bound device
 * current for the calling thread for the duration of the call, so calls from other threads or after cudaSetDevice are safe;
 * a call's non-NULL data pointers must all be host pointers or all be device pointers (EDISON_E_ARG otherwise). */
int edison_init(const edison_cfg *cfg);
void edison_shutdown(void);
const char *edison_last_error(void);             /* text of the last failure in this thread     */
int edison_set_softmax_mode(int mode);
/* number of kernel launches issued by this library since init (bench.py's gpu_launches) */
unsigned long long edison_launch_count(void);
/* pinned host memory for callers that want copies to overlap compute */
void *edison_host_alloc(size_t bytes);
void edison_host_free(void *p);

/* ---- section 1: the firmware's entry points (batch-of-one wrappers) -------------------------- */
/* firmware/src/audioprocessing.h:21-22.  *oup points at a library-owned buffer of 32 q15 values
 * that stays valid until the next call (audioprocessing.c:80, 210).  Not re-entrant, like the
 * original; calls are serialised on one internal lock. */
void audioInit(void);
void audioCalcMFCCs(int16_t *inp, int16_t **oup);

/* firmware/src/ai/ai.h:74-80 */
int aiInitialize(void);
void aiGetInputShape(uint16_t *x, uint16_t *y);
int aiRunInference(void *in_data, void *out_data);   /* in: int8[403], out: int8[10] softmax    */
const char *aiGetKeywordFromIndex(uint32_t idx);
uint32_t aiGetKeywordCount(void);

/* firmware/src/ai/ai_nnom.h:4-9 */
void aiNnomInit(void);
int aiNnomRunInference(void *in_data, void *out_data);
int8_t *aiNnomGetInputBuffer(void);
int8_t *aiNnomGetOutputBuffer(void);

/* firmware/src/app.c:152-153, 675-719 (static there; exported here because the streaming glue
 * of a host application needs them).  They operate on the library-owned 31x13 window that
 * edison_net_input_window() returns. */
void mfccToNetInput(int16_t *mfcc, uint16_t in_x, uint16_t in_y, uint32_t xoffset);
void mfccToNetInputPush(int16_t *mfcc, uint16_t in_x, uint16_t in_y);
int8_t *edison_net_input_window(void);

/* ---- section 2: batched entry points (host OR device pointers, detected per call) ------------ */
/* audioCalcMFCCs over n_frames consecutive 1024-sample frames: mfcc[n_frames][32] q15. */
int edison_mfcc_q15_batch(const int16_t *pcm, size_t n_frames, int16_t *mfcc);

/* appHifMfccAndInference (firmware/src/app.c:167-221) over n_clips clips.  Clip c starts at
 * pcm + c * clip_stride_samples and supplies 31 frames.  Any of the outputs may be NULL:
 * logits[n][10] (Dense output, Q5.2), softmax[n][10] (Q0.7), net_in[n][403]. */
int edison_kws_batch(const int16_t *pcm, size_t n_clips, size_t clip_stride_samples,
                     int8_t *logits, int8_t *softmax, int8_t *net_in);

/* aiRunInference over n precomputed net inputs: net_in[n][403] -> logits[n][10], softmax[n][10] */
int edison_infer_q7_batch(const int8_t *net_in, size_t n, int8_t *logits, int8_t *softmax);

/* Generalised front-end: frames_per_clip frames from each of n_clips clips; writes
 * mfcc[n_clips*frames_per_clip][32] and/or net_in[n_clips*frames_per_clip][13] (either NULL). */
int edison_mfcc_q15_clips(const int16_t *pcm, size_t n_clips, size_t frames_per_clip,
                          size_t clip_stride_samples, int16_t *mfcc, int8_t *net_in);

/* ---- section 2b: the reference's host-side float MFCCs (audio/edison/mfcc/mfcc_utils.py) ----------------------------
 * Fixed to the reference's configuration (audio/config.py:12-36): fs 16 kHz, 1024-sample frames, hop 1024, 32 mel bins
 * 80..7600 Hz.  Output [n][32].  The reference computes in float64 (numpy):
 *   edison_mfcc_f64_*: fp64 on the device, the parity path: max-abs <= 1e-3 / rmse <= 1e-4 on ANY input (observed ~1e-9
 *                      on ordinary audio; the slack is for frames where log(mel + 1e-6) sees only rounding noise);
 *   edison_mfcc_f32_*: fp32 on the device, the fast path: same tolerance on frames whose mel-band energies lie within
 *                      1e4 of each other (recorded audio); synthetic pure tones exceed it. */
#define EDISON_MFCC_F32_HOST     0  /* mfcc():     |FFT| bins 0..511 -> mel -> log(.+1e-6) -> DCT-II / sqrt(64), :134-199 */
#define EDISON_MFCC_F32_MCU      1  /* mfcc_mcu(): fft/1024, |.|/sqrt 2, 513 bins, 128 W then /128, DCT-II / 64, :255-323  */
#define EDISON_MFCC_F32_MCU_LOG  2  /* mfcc_mcu(use_log=True)                                                              */
int edison_mfcc_f64_batch(const int16_t *pcm, size_t n_frames, double *mfcc, int mode);
int edison_mfcc_f64_clips(const int16_t *pcm, size_t n_clips, size_t frames_per_clip, size_t clip_stride_samples,
                          double *mfcc, int mode);
int edison_mfcc_f64_clips_dev(const int16_t *d_pcm, size_t n_clips, size_t frames_per_clip, size_t clip_stride_samples,
                              double *d_mfcc, int mode, void *stream);
int edison_mfcc_f32_batch(const int16_t *pcm, size_t n_frames, float *mfcc, int mode);
int edison_mfcc_f32_clips(const int16_t *pcm, size_t n_clips, size_t frames_per_clip, size_t clip_stride_samples,
                          float *mfcc, int mode);
int edison_mfcc_f32_clips_dev(const int16_t *d_pcm, size_t n_clips, size_t frames_per_clip, size_t clip_stride_samples,
                              float *d_mfcc, int mode, void *stream);

/* ---- section 2b': the firmware's float MFCC in C (firmware/src/audio/mfcc.c; mfcc.h:64-67) -------------------------------
 * mfcc_create(num, offset, frame_len, dec_bits, preemph) / mfcc_compute / mfcc_delete keep their names and signatures
 * (section 1'); these are the batched forms.  Frame f starts at pcm + f * hop_samples and holds frame_len samples
 * (frame_len pads to a power of two <= 1024); out[f][num - offset] q7.  fp32 on the device; agreement with the reference
 * build is exact up to ties of round() (<= 1 LSB) on frames whose bands are not empty. */
typedef struct edison_mfccf edison_mfccf;
int edison_mfccf_create(int num_mfcc_features, int feature_offset, int frame_len, int mfcc_dec_bits, float preemph,
                        edison_mfccf **out);
void edison_mfccf_delete(edison_mfccf *h);
int edison_mfccf_n_out(const edison_mfccf *h);
int edison_mfccf_compute_batch(edison_mfccf *h, const int16_t *pcm, size_t n_frames, size_t hop_samples, int8_t *out);
int edison_mfccf_compute_batch_dev(edison_mfccf *h, const int16_t *d_pcm, size_t n_frames, size_t hop_samples, int8_t *d_out,
                                   void *stream);
/* section 1': firmware/src/audio/mfcc.h:64-67.  mfcc_t is opaque here; mfcc_compute is a batch-of-one wrapper. */
typedef struct _mfcc_t mfcc_t;
mfcc_t *mfcc_create(int num_mfcc_features, int feature_offset, int frame_len, int mfcc_dec_bits, float preemph);
void mfcc_delete(mfcc_t *mfcc);
void mfcc_compute(mfcc_t *mfcc, const int16_t *audio_data, int8_t *mfcc_out);

/* ---- section 2c: the float CNN (Cube.AI build of kws_conv; ai_kws_run via cubeNetRun, firmware/src/ai/ai.c:354-378) ----
 * net_in[n][403] float = MFCC values cast to float, unscaled (firmware/src/app.c:679-683) -> logits[n][10] (dense output)
 * and softmax[n][10] (natural e); either output may be NULL.  fp32 FFMA; tolerance-checked against an fp64 restatement
 * (max-abs 1e-4 on softmax); parity with keras / ST's closed runtime is unpinned. */
int edison_infer_f32_batch(const float *net_in, size_t n, float *logits, float *softmax);
int edison_infer_f32_batch_dev(const float *d_net_in, size_t n, float *d_logits, float *d_softmax, void *stream);

/* ---- section 2d: any NNoM int8 model (nnom_model_create() of a generated weights.h, e.g.
 * firmware/src/ai/nnom/kws_nnom/weights.h:138-161 or the upstream examples under audio/edison/deploy/nnom/examples), converted
 * by tools/convert_nnom_weights.py (version-1 blob: the reference model's conv/dense chain) or tools/convert_nnom_graph.py
 * (version-2 blob: a graph).  Node vocabulary: Input, Conv2D (any kernel and stride, PADDING_VALID / PADDING_SAME), ReLU tails,
 * MaxPool (any kernel, stride, padding), GlobalMaxPool, Concat on the channel axis (equally shaped inputs), Add with output
 * shift, Dense, Softmax.  Refused at load, because the reference's NNoM version does not compute them reproducibly: Concat of
 * unequal inputs (nnom_concat.c:104, :208) and UpSample (nnom_local.c:383).
 * in[n][in_h*in_w*in_c] int8 -> logits[n][n_out] (the softmax layer's input), softmax[n][n_out] (mode of
 * edison_set_softmax_mode); host or device pointers (device pointers 4-byte aligned).  Bit-exact with the NNoM runtime
 * (local backend); generic kernels, not K2. */
typedef struct edison_model edison_model;
int edison_model_load(const char *blob_path, edison_model **out);
void edison_model_free(edison_model *m);
int edison_model_shape(const edison_model *m, int *in_h, int *in_w, int *in_c, int *n_out);
int edison_model_infer(edison_model *m, const int8_t *in, size_t n, int8_t *logits, int8_t *softmax);
/* debug / measurement: naive != 0 runs every node on the one-thread-per-output kernels instead of the tensor-core
 * (mma.sync int8) conv / dense nodes used when a node's input channels are a multiple of 4; results are identical */
int edison_model_set_naive(edison_model *m, int naive);

/* ---- section 3: device-resident variants (all pointers are device pointers; stream is a
 *      cudaStream_t passed as void*, NULL = the library's own non-blocking stream; asynchronous.
 *      Work queued on other streams is NOT ordered against the library's stream: pass the stream the
 *      inputs were produced on, or synchronise first) ------------------------------------------ */
int edison_mfcc_q15_clips_dev(const int16_t *d_pcm, size_t n_clips, size_t frames_per_clip,
                              size_t clip_stride_samples, int16_t *d_mfcc, int8_t *d_net_in,
                              void *stream);
int edison_infer_q7_batch_dev(const int8_t *d_net_in, size_t n, int8_t *d_logits, int8_t *d_softmax,
                              void *stream);
int edison_kws_batch_dev(const int16_t *d_pcm, size_t n_clips, size_t clip_stride_samples,
                         int8_t *d_logits, int8_t *d_softmax, int8_t *d_net_in, void *stream);
/* as the two above, plus a class histogram accumulated by K2's dense epilogue: d_class_counts[10] (int32, device) gets
 * += the number of clips whose FIRST maximum logit is class c.  The caller zeroes it.  Any of d_logits / d_softmax may be
 * NULL (counts only).  This is what a sharded job reduces across GPUs after its last batch. */
int edison_kws_counts_dev(const int16_t *d_pcm, size_t n_clips, size_t clip_stride_samples, int8_t *d_logits,
                          int8_t *d_softmax, int8_t *d_net_in, int32_t *d_class_counts, void *stream);
int edison_infer_q7_counts_dev(const int8_t *d_net_in, size_t n, int8_t *d_logits, int8_t *d_softmax,
                               int32_t *d_class_counts, void *stream);
int edison_sync(void *stream);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it).  edison_kws_batch_dev without the caller-provided
 * d_net_in scratch: the [n_clips][403] network input that K1 writes and K2 reads lives in a static device array of
 * EDISON_NOALLOC_MAX_CLIPS x 403 bytes (3.15 MiB, held by every GPU that loads the library until the process exits).
 * The call performs no cudaMalloc / cudaFree / cudaMallocHost and calls no allocating library routine; it therefore does
 * NOT initialise the library: EDISON_E_NOINIT unless edison_init() succeeded before.  n_clips beyond the maximum is
 * EDISON_E_ARG (split the batch or use edison_kws_batch_dev).  Asynchronous on `stream` like the other _dev variants;
 * results are the same bytes edison_kws_batch_dev produces.  All calls share the one array: calls whose kernels may
 * overlap on the GPU (different streams without a dependency) race; calls on one stream are ordered and safe.  Nothing
 * in the array needs clearing between calls (K1 stores every byte K2 reads).  The first call on a device looks the
 * array's address up (cudaGetSymbolAddress, no allocation): make it once before starting a graph capture. */
#define EDISON_NOALLOC_MAX_CLIPS 8192
int edison_kws_batch_noalloc_dev(const int16_t *d_pcm, size_t n_clips, size_t clip_stride_samples,
                                 int8_t *d_logits, int8_t *d_softmax, void *stream);

/* ---- section 3b: continuous streaming (appMicMfccInfereContinuous, firmware/src/app.c:288-394, and its ISR
 *      appAudioEvent app.c:635-661) over n concurrent streams pinned to this GPU.  Per hop of 1024 new samples
 *      per stream: one MFCC row, mfccToNetInputPush (window of 31 rows advances by one), one full inference,
 *      netOutFilt = 0.9 netOutFilt + (1.0 - 0.9) out (double arithmetic, stored as float; app.c:35-39, 342),
 *      argmax (arm_max_f32: first maximum), "spotted" when the maximum exceeds 0.5 (app.c:34, 344-350).
 *      Like the firmware, inference and the filter start once the window has been filled and one more hop
 *      has arrived (hop count >= 32); before that detect[] is -2 and filt[] stays 0.
 *      On the device a hop is two launches: K1 stores each stream's row straight into its window (a ring of rows),
 *      K2 runs the windows with the filter / argmax / threshold fused into its last epilogue. ------------------ */
typedef struct edison_streams edison_streams;
int edison_streams_open(size_t n_streams, edison_streams **out);
void edison_streams_close(edison_streams *s);
/* hop_pcm[n_streams][1024] host or device.  Outputs (host or device, any may be NULL):
 * softmax[n][10] int8, filt[n][10] float, detect[n] int32 = class index if spotted, -1 if not, -2 while filling.
 * Synchronous: = edison_streams_submit_hop + edison_streams_collect_hop. */
int edison_streams_push_hop(edison_streams *s, const int16_t *hop_pcm, int8_t *softmax, float *filt, int32_t *detect);
/* Pipelined form for host callers: up to two hops may be in flight.  submit starts the host -> device copy of the hop on a copy
 * stream and queues its kernels behind it (hop_pcm must stay valid until the matching collect returns; pinned memory lets the
 * copy of hop h + 1 overlap the kernels of hop h); collect waits for the OLDEST hop in flight and hands out its results. */
int edison_streams_submit_hop(edison_streams *s, const int16_t *hop_pcm);
int edison_streams_collect_hop(edison_streams *s, int8_t *softmax, float *filt, int32_t *detect);
/* device pointers + stream (NULL = the handle's own stream), asynchronous */
int edison_streams_push_hop_dev(edison_streams *s, const int16_t *d_hop_pcm, int8_t *d_softmax, float *d_filt,
                                int32_t *d_detect, void *stream);
unsigned long long edison_streams_hops(const edison_streams *s);
/* 1 (default): K2's incremental kernel -- a window that starts four rows later is the old one moved up by one row of pooled conv2 and of
 *    conv3 output, so a hop computes only the bottom row of every layer (60 of the full window's 153 tensor-core instructions) and takes the
 *    older rows from a row cache in HBM (6.5 KB per stream); bit-identical to the full window, ~20 % faster per hop;
 * 0: the whole 31-row window goes through K2 every hop, as the firmware does (no cache).  Only before the first hop. */
int edison_streams_set_incremental(edison_streams *s, int on);
/* copy one stream's 31x13 window (netInput, app.c:71-93), oldest row first, to the host, for inspection */
int edison_streams_window(edison_streams *s, size_t stream_index, int8_t *window403);

/* ---- section 4: self-checks used by the tests (device arithmetic against itself) ------------- */
/* arm_sqrt_q31 (arm_sqrt_q31.c:50-135) evaluated on the device for n host values */
int edison_debug_sqrt_q31(const int32_t *in, size_t n, int32_t *out);
/* exhaustive check over s in [lo, hi): counts where the fast magnitude path disagrees with the
 * exact Newton path on the device, and a 64-bit checksum of the exact results */
int edison_debug_mag_sweep(uint32_t lo, uint32_t hi, unsigned long long *mismatches,
                           unsigned long long *checksum);
/* K2 with clock64() stamps per schedule unit of block 0's second batch (d_trace: 96 int64 on the device; tools/k2_trace.py) */
int edison_debug_k2_trace_dev(const int8_t *d_net_in, size_t n, int8_t *d_logits, int8_t *d_softmax, long long *d_trace, void *stream);
/* the same stamps from K2 inside the streaming hops that follow (any mode); NULL switches it off */
int edison_streams_debug_trace(edison_streams *s, long long *d_trace);

#ifdef __cplusplus
}
#endif
#endif /* EDISON_B200_H */
