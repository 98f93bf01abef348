#!/usr/bin/env python3
"""Generates tests/golden/mfcc_c_float.npz from the UNMODIFIED reference firmware/src/audio/mfcc.c built for the host
(oracle/_ref/libref_mfcc_f32.so, recipe oracle/build_ref.sh).  Configuration of the reference's only call site,
app.c:540: mfcc_create(13, 1, 512, 8, 0.97f), frames of 512 with hop 256 (app.c:567-584); plus dec_bits = 2 so that the
q7 outputs are not mostly saturated, and frame_len 1024 / 400 (padded to 512)."""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))
from common import synth_frames  # noqa: E402

L = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_mfcc_f32.so"))
L.ref_mfccf_create.restype = ctypes.c_void_p
L.ref_mfccf_create.argtypes = [ctypes.c_int] * 4 + [ctypes.c_float]
L.ref_mfccf_compute.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_void_p]

wav = np.load(os.path.join(HERE, "mfcc_float.npz"))["wav_pcm"]
syn = synth_frames(24, seed=57).reshape(-1)
out = {"wav_pcm": wav, "syn_pcm": syn}
for tag, (num, off, flen, dec, hop) in {"app": (13, 1, 512, 8, 256), "dec2": (13, 1, 512, 2, 256), "f1024": (13, 0, 1024, 1, 1024),
                                        "f400": (10, 2, 400, 3, 160)}.items():
    h = L.ref_mfccf_create(num, off, flen, dec, 0.97)
    for src, pcm in (("wav", wav), ("syn", syn)):
        n = 1 + (pcm.size - flen) // hop
        q = np.zeros((n, num - off), dtype=np.int8)
        L.ref_mfccf_compute(h, pcm.ctypes.data, n, hop, q.ctypes.data)
        out["%s_%s" % (src, tag)] = q
    out["cfg_" + tag] = np.array([num, off, flen, dec, hop])
np.savez_compressed(os.path.join(HERE, "mfcc_c_float.npz"), **out)
print({k: v.shape for k, v in out.items()})
