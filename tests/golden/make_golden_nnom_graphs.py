#!/usr/bin/env python3
"""Generates tests/golden/nnom_graphs.npz and the version-2 blobs tests/golden/nnom_graph_<model>.bin.
Models: upstream NNoM examples shipped inside the reference tree whose graphs go beyond the reference model's chain --
  mnist-simple   (28x28x1: SAME-padded 3x3 convs, 2x2 / stride 2 SAME max pools on even and odd sizes, two dense layers)
  uci-inception  (1x128x9: 1xk convs with stride (1,2) and SAME padding, three parallel branches, Concat, dense + ReLU)
  keyword_spotting (the chain model again, this time through the graph converter)
Outputs come from the reference's UNMODIFIED NNoM runtime (local backend) built for the host by oracle/build_ref.sh
(oracle/_ref/libref_nnom_ex_*.so, libref_nnom_example_local.so).  Blobs are the headers converted by
tools/convert_nnom_graph.py (model data, like a checkpoint)."""
import ctypes
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
EX = "/root/reference/audio/edison/deploy/nnom/examples"
MODELS = [("mnist_simple", "mnist-simple", "libref_nnom_ex_mnist_simple.so", 48),
          ("uci_inception", "uci-inception", "libref_nnom_ex_uci_inception.so", 48),
          ("keyword_spotting", "keyword_spotting", "libref_nnom_example_local.so", 24)]
out = {}
for key, d, lib, n in MODELS:
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "convert_nnom_graph.py"), "%s/%s/mcu/weights.h" % (EX, d),
                           os.path.join(HERE, "nnom_graph_%s.bin" % key)], stdout=subprocess.DEVNULL)
    L = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", lib))
    ni, no = L.ref_example_in_size(), L.ref_example_out_size()
    rng = np.random.default_rng(20)
    x = np.zeros((n, ni), dtype=np.int8)
    for i in range(n):                  # amplitudes from a few LSB to full scale; smooth images too (pooling ties)
        amp = [2, 8, 32, 64, 128][i % 5]
        x[i] = rng.integers(-amp, amp, ni).clip(-128, 127)
        if i % 4 == 3:
            x[i] = (np.cumsum(rng.integers(-3, 4, ni)) % 256 - 128).astype(np.int8)
    x[0] = 0
    x[1] = 127
    x[2] = -128
    lg = np.zeros((n, no), dtype=np.int8)
    sm = np.zeros((n, no), dtype=np.int8)
    assert L.ref_example_run_batch(x.ctypes.data, n, lg.ctypes.data, sm.ctypes.data) == 0
    out[key + "_x"], out[key + "_logits"], out[key + "_softmax_local"] = x, lg, sm
    print(key, ni, no, lg[5], sm[5])
np.savez_compressed(os.path.join(HERE, "nnom_graphs.npz"), **out)
