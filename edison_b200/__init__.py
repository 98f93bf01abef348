"""edison_b200 -- host-side Python mirror of the C ABI in include/edison_b200.h.

The reference's host tooling is Python (audio/edison/**), so tests and bench.py drive
libedison_b200.so through ctypes here.  Nothing in this package computes: every call goes
to the CUDA library and raises if it is missing or no GPU is usable.
"""
from . import capi  # noqa: F401
from . import kws  # noqa: F401
from .capi import (  # noqa: F401
    EdisonError,
    FRAME_LEN,
    N_CLASSES,
    N_COEFF,
    N_FRAMES,
    MFCC_F32_HOST,
    MfccC,
    Model,
    MFCC_F32_MCU,
    MFCC_F32_MCU_LOG,
    N_MEL,
    NET_IN,
    SOFTMAX_CMSIS_REF,
    SOFTMAX_CMSIS_SIMD,
    SOFTMAX_NNOM_LOCAL,
    Streams,
    infer_f32_batch,
    infer_q7_batch,
    init,
    kws_batch,
    launch_count,
    lib,
    lib_path,
    mfcc_float_batch,
    mfcc_float_clips,
    mfcc_q15_batch,
    mfcc_q15_clips,
    set_softmax_mode,
    shutdown,
)
