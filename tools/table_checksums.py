#!/usr/bin/env python3
"""FNV-1a checksums of the CMSIS-DSP q15 tables the front-end needs, computed from the data inside the
reference's prebuilt firmware/src/lib/CMSIS/Lib/GCC/libarm_cortexM4lf_math.a (the C source,
arm_common_tables.c, is missing from the tree: .MISSING_LARGE_BLOBS:4).  libedison_b200 generates
the same tables from formulas at init and refuses to start if its checksums differ from the
constants printed here (edison_b200/csrc/mfcc_q15_kernel.cu: kH1024, kH16, kHAB).

usage: table_checksums.py [/root/reference]
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
from extract_tables import objects  # noqa: E402


def fnv1a(data: bytes) -> int:
    h = 1469598103934665603
    for b in data:
        h ^= b
        h = (h * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    archive = os.path.join(ref, "firmware/src/lib/CMSIS/Lib/GCC/libarm_cortexM4lf_math.a")
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.check_call(["ar", "x", archive, "arm_common_tables.o", "arm_rfft_init_q15.o"], cwd=tmp)
        tabs = {}
        for member in ("arm_common_tables.o", "arm_rfft_init_q15.o"):
            tabs.update(dict(objects(os.path.join(tmp, member))))
    tw1024 = np.frombuffer(tabs["twiddleCoef_1024_q15"], dtype="<i2")
    tw16 = np.frombuffer(tabs["twiddleCoef_16_q15"], dtype="<i2")
    A = np.frombuffer(tabs["realCoefAQ15"], dtype="<i2")
    B = np.frombuffer(tabs["realCoefBQ15"], dtype="<i2")
    # arm_split_rfft_q15 with fftLenReal 32 reads pairs at stride 2*256 q15 (arm_rfft_init_q15.c:213-222)
    ab = np.concatenate([np.stack([A[512 * n: 512 * n + 2] for n in range(16)]).reshape(-1),
                         np.stack([B[512 * n: 512 * n + 2] for n in range(16)]).reshape(-1)]).astype("<i2")
    print("kH1024 = 0x%016x" % fnv1a(tw1024[:1536].tobytes()))
    print("kH16   = 0x%016x" % fnv1a(tw16[:24].tobytes()))
    print("kHAB   = 0x%016x" % fnv1a(ab.tobytes()))


if __name__ == "__main__":
    main()
