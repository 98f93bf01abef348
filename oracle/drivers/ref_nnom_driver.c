/* Driver around the UNMODIFIED reference NNoM runtime + kws_nnom/weights.h
 * (TEST INFRASTRUCTURE; linked into oracle/_ref/libref_nnom_*.so).
 * Mirrors firmware/src/ai/ai_nnom.c:63-85 (aiNnomInit / aiNnomRunInference) and taps the
 * Dense output (the int8 logits) through model_set_callback (nnom.c:1043). */
#include <stdint.h>
#include <string.h>
#include <stdio.h>
#include <unistd.h>
#include <fcntl.h>
#include "nnom.h"
#include "nnom_local.h"
#include "kws_nnom/weights.h"
#ifdef NNOM_USING_CMSIS_NN
#include "arm_nnfunctions.h"
#endif

static nnom_model_t *model;
static int8_t tap_dense[10];
static int8_t tap_layers[8][4096];   /* conv1, pool1, conv2, pool2, conv3, conv4 (post-activation), dense, softmax */
static int    tap_count;

static nnom_status_t tap_cb(nnom_model_t *m, nnom_layer_t *layer)
{
  (void)m;
  size_t n = tensor_size(layer->out->tensor);
  if (layer->type == NNOM_DENSE) memcpy(tap_dense, layer->out->tensor->p_data, 10);
  if (layer->type != NNOM_INPUT && layer->type != NNOM_OUTPUT && tap_count < 8) {
    if (n > sizeof tap_layers[0]) n = sizeof tap_layers[0];
    memcpy(tap_layers[tap_count++], layer->out->tensor->p_data, n);
  }
  return NN_SUCCESS;
}

int ref_nnom_init(void)
{
  if (model) return 0;
  /* model_compile prints its layer table through NNOM_LOG == printf; keep test output clean */
  fflush(stdout);
  int keep = dup(1), nul = open("/dev/null", O_WRONLY);
  if (nul >= 0) dup2(nul, 1);
  model = nnom_model_create();
  fflush(stdout);
  if (keep >= 0) { dup2(keep, 1); close(keep); }
  if (nul >= 0) close(nul);
  if (!model) return -1;
  model_set_callback(model, tap_cb);
  return 0;
}

/* in: [n][403] int8 ; logits/softmax: [n][10] int8 (either may be NULL) */
int ref_nnom_run_batch(const int8_t *in, long n, int8_t *logits, int8_t *softmax)
{
  if (ref_nnom_init()) return -1;
  for (long i = 0; i < n; i++) {
    memcpy(nnom_input_data, in + i * 403, 403);
    tap_count = 0;
    nnom_status_t r = model_run(model);
    if (r != NN_SUCCESS) return (int)r;
    if (logits) memcpy(logits + i * 10, tap_dense, 10);
    if (softmax) memcpy(softmax + i * 10, nnom_output_data, 10);
  }
  return 0;
}

/* one clip with every layer output: buffer of 8 x 4096 bytes */
int ref_nnom_run_layers(const int8_t *in, int8_t *layers8x4096, int *n_layers)
{
  if (ref_nnom_init()) return -1;
  memcpy(nnom_input_data, in, 403);
  tap_count = 0;
  nnom_status_t r = model_run(model);
  memcpy(layers8x4096, tap_layers, sizeof tap_layers);
  *n_layers = tap_count;
  return (int)r;
}

/* the softmax this build's softmax_run dispatches to (nnom_softmax.c:57-66) */
void ref_softmax_q7(const int8_t *in, int n, int8_t *out)
{
#ifdef NNOM_USING_CMSIS_NN
  arm_softmax_q7(in, (uint16_t)n, out);
#else
  local_softmax_q7(in, (uint32_t)n, out);
#endif
}
