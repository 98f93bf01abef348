/* Host shim for firmware/src/io/hostinterface.h (TEST INFRASTRUCTURE).
 * hiSendS16 is recorded by the driver so per-stage dumps (audioDumpToHost) are observable. */
#pragma once
#include <stdint.h>
typedef enum { DATA_FORMAT_S16 = '3' } hiDataFormat_t;
void hiSendS16(int16_t *data, uint32_t len, uint8_t tag);
uint32_t hiReceive(void *data, uint32_t maxlen, hiDataFormat_t fmt, uint8_t *tag);
