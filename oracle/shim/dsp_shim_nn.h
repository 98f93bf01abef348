/* Force-included AFTER arm_math.h for the CMSIS-NN "SIMD" host build (TEST INFRASTRUCTURE).
 * Four Cortex-M4 intrinsics used by CMSIS-NN are not emulated by arm_math.h's host
 * mode; their architectural semantics (ARMv7-M ARM, A7.7) are restated here. */
#pragma once
#include <stdint.h>
static inline uint32_t __ROR(uint32_t v, uint32_t s) { s &= 31u; return s ? ((v >> s) | (v << (32u - s))) : v; }
/* per-byte modular signed add / subtract (no saturation) */
static inline uint32_t __SADD8(uint32_t x, uint32_t y)
{
  uint32_t r = 0;
  for (int i = 0; i < 4; i++) r |= (uint32_t)(uint8_t)((uint8_t)(x >> (8 * i)) + (uint8_t)(y >> (8 * i))) << (8 * i);
  return r;
}
static inline uint32_t __SSUB8(uint32_t x, uint32_t y)
{
  uint32_t r = 0;
  for (int i = 0; i < 4; i++) r |= (uint32_t)(uint8_t)((uint8_t)(x >> (8 * i)) - (uint8_t)(y >> (8 * i))) << (8 * i);
  return r;
}
/* per-halfword unsigned saturate of signed halfwords to [0, 2^sat - 1] */
static inline uint32_t __USAT16(uint32_t x, uint32_t sat)
{
  int32_t lo = (int16_t)(x & 0xFFFFu), hi = (int16_t)(x >> 16);
  int32_t mx = (int32_t)((1u << sat) - 1u);
  lo = lo < 0 ? 0 : (lo > mx ? mx : lo);
  hi = hi < 0 ? 0 : (hi > mx ? mx : hi);
  return ((uint32_t)hi << 16) | (uint32_t)lo;
}
#define ARM_MATH_DSP 1
