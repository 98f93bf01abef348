// nnom_q7_kernel.cu -- the model loader of K2, the int8 NNoM keyword-spotting net (firmware/src/ai/nnom/kws_nnom/weights.h:138-161).
//
// The kernels are in nnom_q7_umma.cu (tcgen05.mma.kind::i8, TMEM accumulators): the whole 31 x 13 window through the net, and the
// incremental sliding-window variant of the streaming path (one new row per layer and hop, older rows from a row cache in HBM).
// This file turns the blob made by tools/convert_nnom_weights.py into the weight image those kernels read (k2_umma_plan.h
// build_image) and owns it: the opaque NnomModel handle of edison_internal.h IS the device pointer of that image.
//
// Arithmetic reproduced: arm_convolve_HWC_q7_{basic,fast}_nonsquare / local_convolve_HWC_q7_nonsquare (nnom_local.c:425-484),
// local_maxpool_q7_HWC (nnom_local.c:117-160), arm_fully_connected_q7_opt (weights de-interleaved at conversion time) and the
// three softmax variants:   out = SSAT8( ((bias << bias_lshift) + (1 << (out_rshift-1)) + sum in*w) >> out_rshift )
// int32 accumulation is exact and order independent (|acc| < 2^24), so any MMA tiling gives the reference's bits.
// Round 1's mma.sync kernels (full window and incremental) are gone: there is no legacy tensor path behind K2.
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "edison_internal.h"
#include "../../include/edison_b200.h"
#include "nnom_blob.h"

namespace edison {

cudaError_t nnom_configure(void) { return nnom_umma_configure(); }

cudaError_t launch_nnom_q7(const NnomModel *model, const K2Launch &kl, int sm_count, cudaStream_t stream) {
  if (!model) return cudaErrorNotReady;
  return launch_nnom_q7_umma(reinterpret_cast<const uint8_t *>(model), kl, sm_count, stream);
}

// blob (tools/convert_nnom_weights.py) -> weight image on the device
int nnom_load(const char *path, NnomModel **out_dev, int *in_scale, int *in_min, int *in_max, char *err, size_t errlen) {
  KwsBlob blob;
  int rc = read_kws_blob(path, blob, err, errlen);
  if (rc) return rc;
  *in_scale = blob.in_scale; *in_min = blob.in_min; *in_max = blob.in_max;
  const int8_t *w[5], *b[5];
  int rs[5], ls[5];
  for (int i = 0; i < 5; i++) { w[i] = blob.L[i].w; b[i] = blob.L[i].b; rs[i] = (int)blob.L[i].h[5]; ls[i] = (int)blob.L[i].h[6]; }
  uint8_t *image = nullptr;
  rc = nnom_umma_upload(w, b, rs, ls, &image);
  if (rc) {
    cudaGetLastError();
    snprintf(err, errlen, "upload of the K2 weight image failed");
    return rc;
  }
  *out_dev = reinterpret_cast<NnomModel *>(image);
  return EDISON_OK;
}

void nnom_free(NnomModel *m) {
  if (m) cudaFree(m);
}

}  // namespace edison
