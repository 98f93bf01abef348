#!/usr/bin/env python3
"""Generates tests/golden/mfcc_float.npz by running the REFERENCE's own Python (audio/edison/mfcc/mfcc_utils.py) in
the build container, where /root/reference is mounted.  The fixture travels to the GPU box; the reference does not.
Inputs: the reference's sample WAV (audio/data/edison_16k_16b.wav, 11,243 int16 samples -> 10 frames, BASELINE config #1)
and seeded synthetic frames (speech-like, full-scale uniform, low-level uniform)."""
import os
import sys
import wave

import numpy as np

REF = "/root/reference/audio"
sys.path.insert(0, REF)
import edison.mfcc.mfcc_utils as mfu  # noqa: E402  (do not import edison.mfcc.mfcc: it pulls matplotlib)
from config import fs, frame_length, frame_step, num_mel_bins, lower_edge_hertz, upper_edge_hertz, mel_mtx_scale  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from common import synth_frames  # noqa: E402


def run(fn, data, *extra):
    o = fn(data, fs, data.shape[0], frame_length, frame_step, 0, frame_length, num_mel_bins, lower_edge_hertz,
           upper_edge_hertz, *extra)
    return np.array([f["mfcc"] for f in o]), np.array([f["mel_spectrogram"] for f in o])


with wave.open(os.path.join(REF, "data", "edison_16k_16b.wav")) as w:
    assert w.getframerate() == 16000 and w.getsampwidth() == 2 and w.getnchannels() == 1
    wav = np.frombuffer(w.readframes(w.getnframes()), dtype=np.int16).copy()
syn = synth_frames(48, seed=31).reshape(-1)
out = {"wav_pcm": wav, "syn_pcm": syn}
for name, pcm in (("wav", wav), ("syn", syn)):
    out[name + "_host_mfcc"], out[name + "_host_mel"] = run(mfu.mfcc, pcm)
    out[name + "_mcu_mfcc"], out[name + "_mcu_mel"] = run(mfu.mfcc_mcu, pcm, mel_mtx_scale)
    out[name + "_mculog_mfcc"], _ = run(mfu.mfcc_mcu, pcm, mel_mtx_scale, True)
out["mel_w512"] = mfu.gen_mel_weight_matrix(num_mel_bins, 512, fs, lower_edge_hertz, upper_edge_hertz)
out["mel_w513"] = mfu.gen_mel_weight_matrix(num_mel_bins, 513, fs, lower_edge_hertz, upper_edge_hertz)
np.savez_compressed(os.path.join(HERE, "mfcc_float.npz"), **out)
print("wav frames", out["wav_host_mfcc"].shape, "mfcc[0][:4]", out["wav_host_mfcc"][0][:4], "mfcc[3][:4]", out["wav_host_mfcc"][3][:4])
print("mcu[3][:4]", out["wav_mcu_mfcc"][3][:4])
