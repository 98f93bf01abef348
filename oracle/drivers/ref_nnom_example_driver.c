/* TEST INFRASTRUCTURE.  Driver around the UNMODIFIED reference NNoM runtime running the upstream keyword_spotting example
 * model (audio/edison/deploy/nnom/examples/keyword_spotting/mcu/weights.h: Input 63x12x1, 35 classes) -- a second real
 * NNoM model of the same layer family, used to pin the generic executor (edison_model_*) and the blob converter. */
#include <stdint.h>
#include <string.h>
#include <stdio.h>
#include <unistd.h>
#include <fcntl.h>
#include "nnom.h"
#include "nnom_local.h"
#include "weights.h"

static nnom_model_t *model;
static int8_t tap_dense[64];

static nnom_status_t tap_cb(nnom_model_t *m, nnom_layer_t *layer)
{
  (void)m;
  if (layer->type == NNOM_DENSE) memcpy(tap_dense, layer->out->tensor->p_data, sizeof nnom_output_data);
  return NN_SUCCESS;
}

int ref_example_in_size(void) { return (int)sizeof nnom_input_data; }
int ref_example_out_size(void) { return (int)sizeof nnom_output_data; }

int ref_example_run_batch(const int8_t *in, long n, int8_t *logits, int8_t *softmax)
{
  if (!model) {
    fflush(stdout);
    int keep = dup(1), nul = open("/dev/null", O_WRONLY);
    if (nul >= 0) dup2(nul, 1);
    model = nnom_model_create();
    fflush(stdout);
    if (keep >= 0) { dup2(keep, 1); close(keep); }
    if (nul >= 0) close(nul);
    if (!model) return -1;
    model_set_callback(model, tap_cb);
  }
  const long ni = (long)sizeof nnom_input_data, no = (long)sizeof nnom_output_data;
  for (long i = 0; i < n; i++) {
    memcpy(nnom_input_data, in + i * ni, (size_t)ni);
    nnom_status_t r = model_run(model);
    if (r != NN_SUCCESS) return (int)r;
    if (logits) memcpy(logits + i * no, tap_dense, (size_t)no);
    if (softmax) memcpy(softmax + i * no, nnom_output_data, (size_t)no);
  }
  return 0;
}
