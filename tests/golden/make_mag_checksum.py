#!/usr/bin/env python3
"""The 64-bit checksum that edison_debug_mag_sweep(0, 0x80000001) returns, computed from the REFERENCE's own arm_sqrt_q31
(oracle/_ref/libref_mfcc_simd.so, built from the unmodified firmware/src/lib/CMSIS/DSP/Source/FastMathFunctions/arm_sqrt_q31.c):
    sum over s in [0, 2^31] of  (uint32)(int16)(arm_sqrt_q31(s) >> 16) * ((s & 1023) + 1)   (mod 2^64)
i.e. every value cmpl_mag_q15 (firmware/src/audioprocessing.c:294-307) can produce, weighted so that a shifted or swapped result
changes the sum.  Run in the build container (all cores, about a minute); prints the constant that tests/test_gpu_direct.py holds.
    python tests/golden/make_mag_checksum.py
"""
import ctypes
import multiprocessing as mp
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CHUNK = 1 << 24


def part(k):
    L = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_mfcc_simd.so"))
    lo = k * CHUNK
    hi = min(lo + CHUNK, (1 << 31) + 1)
    s = np.arange(lo, hi, dtype=np.uint64)
    x = s.astype(np.uint32).view(np.int32)           # 2^31 wraps to INT_MIN: arm_sqrt_q31 returns 0 for it, like the device path
    out = np.zeros(x.size, np.int32)
    L.ref_sqrt_q31_batch(x.ctypes.data_as(ctypes.c_void_p), ctypes.c_long(x.size), out.ctypes.data_as(ctypes.c_void_p))
    mag = (out >> 16).astype(np.int16).astype(np.int32).view(np.uint32).astype(np.uint64)
    w = (s & np.uint64(1023)) + np.uint64(1)
    return int((mag * w).sum(dtype=np.uint64))       # wraps mod 2^64 like the device accumulator


def main():
    n = ((1 << 31) + 1 + CHUNK - 1) // CHUNK
    with mp.Pool(os.cpu_count()) as pool:
        total = sum(pool.map(part, range(n))) & ((1 << 64) - 1)
    print("MAG_SWEEP_CHECKSUM = 0x%016x" % total)


if __name__ == "__main__":
    main()
