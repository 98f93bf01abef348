/* TEST INFRASTRUCTURE.  Thin host driver around the reference's float MFCC (firmware/src/audio/mfcc.c, compiled unmodified
 * where it lies).  mfcc.c sizes two of its allocations for a 32-bit target (pointer array as sizeof(float)*26, :127; a
 * precedence slip in :163; and mfcc_malloc clears sizeof(mfcc_t) bytes whatever was asked, :34-41), which overruns on a
 * 64-bit host; build_ref.sh therefore compiles mfcc.c with -Dmalloc=ref_pad_malloc, and this file supplies an allocator
 * that hands out generously padded zeroed blocks.  The arithmetic is untouched. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "mfcc.h"

void *ref_pad_malloc(size_t size) { return calloc(1, 4 * size + 1024); }

void *ref_mfccf_create(int num_mfcc_features, int feature_offset, int frame_len, int mfcc_dec_bits, float preempha) {
  return mfcc_create(num_mfcc_features, feature_offset, frame_len, mfcc_dec_bits, preempha);
}

/* n_frames frames of frame_len samples, frame f starts at pcm + f*hop; out[f][num - offset] */
void ref_mfccf_compute(void *h, const int16_t *pcm, size_t n_frames, size_t hop, int8_t *out) {
  mfcc_t *m = (mfcc_t *)h;
  const int n_out = m->num_mfcc_features - m->num_features_offset;
  for (size_t f = 0; f < n_frames; f++) mfcc_compute(m, pcm + f * hop, (q7_t *)out + f * n_out);
}

/* the intermediates of the last frame, for stage-wise comparison: log mel energies [26] */
void ref_mfccf_last_logmel(void *h, float *out26) { memcpy(out26, ((mfcc_t *)h)->mel_energies, 26 * sizeof(float)); }
