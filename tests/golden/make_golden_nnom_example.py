#!/usr/bin/env python3
"""Generates tests/golden/nnom_example.npz and tests/golden/nnom_kws_example.bin.
The model is the upstream NNoM keyword_spotting example shipped inside the reference
(audio/edison/deploy/nnom/examples/keyword_spotting/mcu/weights.h: Input 63x12x1 -> ... -> Dense 35 -> Softmax); the outputs
come from the reference's UNMODIFIED NNoM runtime (local backend) built for the host: oracle/_ref/libref_nnom_example_local.so.
The blob is that header converted by tools/convert_nnom_weights.py (model data, like a checkpoint)."""
import ctypes
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
HDR = "/root/reference/audio/edison/deploy/nnom/examples/keyword_spotting/mcu/weights.h"
subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "convert_nnom_weights.py"), HDR, os.path.join(HERE, "nnom_kws_example.bin")])
L = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_nnom_example_local.so"))
ni, no = L.ref_example_in_size(), L.ref_example_out_size()
rng = np.random.default_rng(20)
n = 96
x = np.zeros((n, ni), dtype=np.int8)
for i in range(n):                      # amplitudes from a few LSB to full scale, plus all-zero and constant inputs
    amp = [2, 8, 32, 64, 128][i % 5]
    x[i] = rng.integers(-amp, amp, ni).clip(-128, 127)
x[0] = 0
x[1] = 127
x[2] = -128
lg = np.zeros((n, no), dtype=np.int8)
sm = np.zeros((n, no), dtype=np.int8)
assert L.ref_example_run_batch(x.ctypes.data, n, lg.ctypes.data, sm.ctypes.data) == 0
np.savez_compressed(os.path.join(HERE, "nnom_example.npz"), x=x, logits=lg, softmax_local=sm)
print(ni, no, lg[5], sm[5])
