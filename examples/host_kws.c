/* host_kws.c -- a C host program written against the FIRMWARE's entry points (audioprocessing.h:21-22, ai.h:74-80,
 * app.c:152-153), linked against libedison_b200.so instead of audioprocessing.c + the NNoM runtime + CMSIS.
 * It is appHifMfccAndInference (firmware/src/app.c:167-221) with the UART transfers replaced by a file:
 *
 *     host_kws <pcm.s16le>        reads 31 * 1024 int16 samples, prints the 403 net-input bytes and the 10 softmax bytes
 *
 * and then the same clip through the batched call, to show that both give the same bytes.
 * Build: gcc -std=c99 -Iinclude examples/host_kws.c -Ledison_b200 -ledison_b200 -Wl,-rpath,$PWD/edison_b200 -o host_kws */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "edison_b200.h"

int main(int argc, char **argv) {
  static int16_t pcm[EDISON_CLIP_SAMPLES];
  if (argc < 2) { fprintf(stderr, "usage: %s <pcm.s16le>\n", argv[0]); return 2; }
  FILE *f = fopen(argv[1], "rb");
  if (!f || fread(pcm, sizeof(int16_t), EDISON_CLIP_SAMPLES, f) != EDISON_CLIP_SAMPLES) { fprintf(stderr, "short read\n"); return 2; }
  fclose(f);

  /* --- the firmware's loop, unchanged ------------------------------------------------------------------- */
  uint16_t in_x, in_y;
  int16_t *out_mfccs;
  int8_t net_out[EDISON_N_CLASSES];
  audioInit();
  if (aiInitialize() != 0) { fprintf(stderr, "aiInitialize: %s\n", edison_last_error()); return 1; }
  aiGetInputShape(&in_x, &in_y);
  for (int frame = 0; frame < in_y; frame++) {
    audioCalcMFCCs(&pcm[frame * EDISON_FRAME_LEN], &out_mfccs);
    mfccToNetInput(out_mfccs, in_x, in_y, (uint32_t)frame);
  }
  int8_t *net_in = edison_net_input_window();
  if (aiRunInference(net_in, net_out) != 0) { fprintf(stderr, "aiRunInference failed\n"); return 1; }
  printf("net_in");
  for (int i = 0; i < EDISON_NET_IN; i++) printf(" %d", net_in[i]);
  printf("\nnet_out");
  for (int i = 0; i < EDISON_N_CLASSES; i++) printf(" %d", net_out[i]);
  int best = 0;
  for (int i = 1; i < EDISON_N_CLASSES; i++) if (net_out[i] > net_out[best]) best = i;
  printf("\nkeyword %s (%u classes)\n", aiGetKeywordFromIndex((uint32_t)best), aiGetKeywordCount());

  /* --- the batched entry point on the same clip ---------------------------------------------------------- */
  int8_t logits[EDISON_N_CLASSES], softmax[EDISON_N_CLASSES], net_in2[EDISON_NET_IN];
  if (edison_kws_batch(pcm, 1, EDISON_CLIP_SAMPLES, logits, softmax, net_in2)) { fprintf(stderr, "%s\n", edison_last_error()); return 1; }
  printf("batched %s\n", memcmp(softmax, net_out, sizeof softmax) == 0 && memcmp(net_in2, net_in, sizeof net_in2) == 0 ? "identical" : "DIFFERENT");
  edison_shutdown();
  return 0;
}
