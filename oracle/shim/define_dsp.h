/* Force-included AFTER arm_math.h in the "firmware-true" host build.
 * arm_math.h's __GNUC_PYTHON__ mode defines C emulations of the Cortex-M4 DSP
 * intrinsics only while ARM_MATH_DSP is unset; the SIMD code branches of the
 * CMSIS sources compile only while it is set (the firmware Makefile passes
 * -DARM_MATH_CM4 -DARM_MATH_DSP).  Defining it here, after the header, selects
 * the SIMD branches on top of the emulated intrinsics. */
#pragma once
#define ARM_MATH_DSP 1
