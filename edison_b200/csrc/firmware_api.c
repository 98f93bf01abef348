/* firmware_api.c -- the firmware's own entry points (include/edison_b200.h section 1) as
 * batch-of-one wrappers over the batched CUDA path.  Plain C, like the host code it replaces:
 *   audioInit / audioCalcMFCCs            firmware/src/audioprocessing.c:99-108, 116-215
 *   aiInitialize .. aiGetKeywordCount     firmware/src/ai/ai.c:205-251
 *   aiNnom*                               firmware/src/ai/ai_nnom.c:63-95
 *   mfccToNetInput / mfccToNetInputPush   firmware/src/app.c:675-719
 * Semantics kept: library-owned static output buffers, not re-entrant, 0 == NN_SUCCESS.
 * A failure of the GPU path cannot be reported through audioCalcMFCCs' void signature; like the
 * firmware's Error_Handler() (audioprocessing.c:198-201) it is fatal: the message goes to stderr
 * and the process aborts rather than return stale data. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/edison_b200.h"

int edison_internal_keyword(uint32_t idx, const char **out);
int edison_internal_keyword_count(void);
int edison_internal_input_quant(int *scale, int *mn, int *mx);

static int16_t bufDctInline[EDISON_N_MEL];        /* audioprocessing.c:80 */
static int8_t nnom_input_data[EDISON_NET_IN];     /* weights.h:136 */
static int8_t nnom_output_data[EDISON_N_CLASSES]; /* weights.h:137 */
static int8_t netInput[EDISON_NET_IN];            /* app.c:71-93 */

static void fatal(const char *who, int rc) {
  fprintf(stderr, "libedison_b200: %s failed (%d): %s\n", who, rc, edison_last_error());
  abort();
}

void audioInit(void) {
  int rc = edison_init(NULL);
  if (rc) fatal("audioInit", rc);
}

void audioCalcMFCCs(int16_t *inp, int16_t **oup) {
  int rc = edison_mfcc_q15_batch(inp, 1, bufDctInline);
  if (rc) fatal("audioCalcMFCCs", rc);
  *oup = bufDctInline;
}

int aiInitialize(void) { return edison_init(NULL); }

void aiGetInputShape(uint16_t *x, uint16_t *y) {
  *x = EDISON_N_COEFF;
  *y = EDISON_N_FRAMES;
}

void aiNnomInit(void) {
  int rc = edison_init(NULL);
  if (rc) fatal("aiNnomInit", rc);
}

int aiNnomRunInference(void *in_data, void *out_data) {
  int rc;
  if (!in_data || !out_data) return EDISON_E_ARG;
  memcpy(nnom_input_data, in_data, sizeof nnom_input_data);
  rc = edison_infer_q7_batch(nnom_input_data, 1, NULL, nnom_output_data);
  if (rc) return rc;
  memcpy(out_data, nnom_output_data, sizeof nnom_output_data);
  return 0;
}

int aiRunInference(void *in_data, void *out_data) { return aiNnomRunInference(in_data, out_data); }

int8_t *aiNnomGetInputBuffer(void) { return nnom_input_data; }
int8_t *aiNnomGetOutputBuffer(void) { return nnom_output_data; }

const char *aiGetKeywordFromIndex(uint32_t idx) {
  const char *s = NULL;
  if (edison_internal_keyword(idx, &s)) return NULL;
  return s;
}

uint32_t aiGetKeywordCount(void) { return (uint32_t)edison_internal_keyword_count(); }

int8_t *edison_net_input_window(void) { return netInput; }

void mfccToNetInput(int16_t *mfcc, uint16_t in_x, uint16_t in_y, uint32_t xoffset) {
  int scale = 1, mn = -128, mx = 127, c;
  (void)in_y;
  if (edison_internal_input_quant(&scale, &mn, &mx)) fatal("mfccToNetInput", EDISON_E_NOINIT);
  for (c = 0; c < in_x; c++) {
    size_t at = (size_t)xoffset * in_x + (size_t)c;
    int16_t t;
    if (at >= sizeof netInput) break;
    t = (int16_t)(mfcc[c] / scale);
    t = (t > mx) ? (int16_t)mx : t;
    t = (t < mn) ? (int16_t)mn : t;
    netInput[at] = (int8_t)t;
  }
}

void mfccToNetInputPush(int16_t *mfcc, uint16_t in_x, uint16_t in_y) {
  size_t keep;
  if (in_y == 0) return;
  keep = (size_t)(in_y - 1) * in_x;
  if (keep + in_x > sizeof netInput) return;
  memmove(netInput, netInput + in_x, keep);
  mfccToNetInput(mfcc, in_x, in_y, (uint32_t)(in_y - 1));
}

/* ---- firmware/src/audio/mfcc.h:64-67: float MFCC handle API (mfcc_t is opaque: the state lives on the device) -------------- */
mfcc_t *mfcc_create(int num_mfcc_features, int feature_offset, int frame_len, int mfcc_dec_bits, float preemph) {
  edison_mfccf *h = NULL;
  if (edison_mfccf_create(num_mfcc_features, feature_offset, frame_len, mfcc_dec_bits, preemph, &h)) return NULL;  /* mfcc_malloc -> NULL */
  return (mfcc_t *)h;
}

void mfcc_delete(mfcc_t *mfcc) { edison_mfccf_delete((edison_mfccf *)mfcc); }

void mfcc_compute(mfcc_t *mfcc, const int16_t *audio_data, int8_t *mfcc_out) {
  int rc = edison_mfccf_compute_batch((edison_mfccf *)mfcc, audio_data, 1, 1, mfcc_out);
  if (rc) fatal("mfcc_compute", rc);
}
