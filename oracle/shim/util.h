/* Host shim for firmware/src/util/util.h (TEST INFRASTRUCTURE). */
#pragma once
#include <stdint.h>
uint8_t utilTic(void);
uint32_t utilToc(uint8_t id);
