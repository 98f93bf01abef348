/* Driver + stubs around the UNMODIFIED reference firmware/src/audioprocessing.c
 * (TEST INFRASTRUCTURE; linked into oracle/_ref/libref_mfcc_{simd,c}.so).
 * Nothing here is part of the product path.  It exposes the reference's
 * audioCalcMFCCs (audioprocessing.c:116) over flat batches and records the
 * per-stage dumps that audioDumpToHost (audioprocessing.c:221-231) sends
 * through hiSendS16 (tags 0 = fft, 1 = spectrum, 2 = mel, 4 = dct). */
#include <stdint.h>
#include <string.h>
#include <stdio.h>
#include <stdlib.h>
#include "main.h"
#include "hostinterface.h"
#include "util.h"
#include "cyclecounter.h"

void audioInit(void);
void audioCalcMFCCs(int16_t *inp, int16_t **oup);
void audioDumpToHost(void);

UART_HandleTypeDef huart1, huart4;
void Error_Handler(void) { fprintf(stderr, "reference Error_Handler()\n"); abort(); }
void mainSetPrintfUart(UART_HandleTypeDef *p) { (void)p; }
uint8_t utilTic(void) { return 0; }
uint32_t utilToc(uint8_t id) { (void)id; return 0; }
void cycProfStart(const char *n) { (void)n; }
void cycProfEvent(const char *e) { (void)e; }
void cycProfStop(void) {}
uint32_t hiReceive(void *data, uint32_t maxlen, hiDataFormat_t fmt, uint8_t *tag)
{ (void)data; (void)maxlen; (void)fmt; (void)tag; return 0; }

static int16_t *rec_dst[8];
static uint32_t rec_max[8];
void hiSendS16(int16_t *data, uint32_t len, uint8_t tag)
{
  if (tag < 8 && rec_dst[tag]) memcpy(rec_dst[tag], data, sizeof(int16_t) * (len < rec_max[tag] ? len : rec_max[tag]));
}

/* out: [n_frames][32] q15 */
void ref_mfcc_batch(const int16_t *pcm, long n_frames, int16_t *out)
{
  int16_t frame[1024];
  int16_t *res;
  audioInit();
  for (long f = 0; f < n_frames; f++) {
    memcpy(frame, pcm + f * 1024, sizeof frame);
    audioCalcMFCCs(frame, &res);
    memcpy(out + f * 32, res, 32 * sizeof(int16_t));
  }
}

/* one frame with all intermediate stages */
void ref_mfcc_stages(const int16_t *pcm, int16_t *fft2048, int16_t *spect1024, int16_t *mel32, int16_t *dct32)
{
  int16_t frame[1024];
  int16_t *res;
  memcpy(frame, pcm, sizeof frame);
  audioInit();
  audioCalcMFCCs(frame, &res);
  rec_dst[0] = fft2048;  rec_max[0] = 2048;
  rec_dst[1] = spect1024; rec_max[1] = 1024;
  rec_dst[2] = mel32;    rec_max[2] = 32;
  rec_dst[4] = dct32;    rec_max[4] = 32;
  audioDumpToHost();
  memset(rec_dst, 0, sizeof rec_dst);
}

/* arm_sqrt_q31 exposed for exhaustive checks of the restatement */
int arm_sqrt_q31(int32_t in, int32_t *pOut);
void ref_sqrt_q31_batch(const int32_t *in, long n, int32_t *out)
{
  for (long i = 0; i < n; i++) arm_sqrt_q31(in[i], &out[i]);
}
