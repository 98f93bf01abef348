#!/usr/bin/env python3
"""Generate tests/golden/*.npz by RUNNING THE REFERENCE'S OWN C (oracle/_ref, built by oracle/build_ref.sh
from the unmodified sources under /root/reference) and its numpy MFCC.  Run in the build container:

    python tests/golden/make_golden.py

Outputs (small, committed):
  q15_mfcc.npz   inputs (probe frames, the two reference WAVs framed, adversarial frames) and the
                 firmware-true (SIMD build) outputs incl. per-stage dumps; plus the plain-C build's outputs
  nnom_q7.npz    seeded int8 net inputs and logits / softmax of the three NNoM builds
"""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import common  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref")
P = lambda a: a.ctypes.data_as(ctypes.c_void_p)


def read_wav_int16(path):
    import scipy.io.wavfile as wf
    fs, x = wf.read(path)
    assert fs == 16000 and x.dtype == np.int16
    return x


def main():
    wav1 = read_wav_int16("/root/reference/audio/data/edison_16k_16b.wav")
    wav2 = read_wav_int16("/root/reference/audio/data/hey_short_16k.wav")
    # hey_short padded to one frame the way mfcc_on_mcu.py:321-323 does (constant_values=(4, 6))
    hey = np.pad(wav2.astype(np.int64), (0, 1024 - wav2.size), constant_values=(4, 6)).astype(np.int16)
    n1 = 1 + (wav1.size - 1024) // 1024
    frames = np.concatenate([common.lcg_probe_frames(), wav1[: n1 * 1024].reshape(n1, 1024), hey[None, :],
                             common.synth_frames(96, seed=20)])
    n = frames.shape[0]
    out = {"pcm": frames}
    for name in ("simd", "c"):
        L = ctypes.CDLL(os.path.join(REF, "libref_mfcc_%s.so" % name))
        mf = np.zeros((n, 32), np.int16)
        L.ref_mfcc_batch(P(frames), ctypes.c_long(n), P(mf))
        out["mfcc_" + name] = mf
        if name == "simd":
            fft = np.zeros((n, 2048), np.int16); spect = np.zeros((n, 1024), np.int16)
            mel = np.zeros((n, 32), np.int16); dct = np.zeros((n, 32), np.int16)
            for i in range(n):
                L.ref_mfcc_stages(P(frames[i]), P(fft[i]), P(spect[i]), P(mel[i]), P(dct[i]))
            assert (dct == mf).all()
            out.update(fft_simd=fft, spect_simd=spect[:, :512], mel_simd=mel)
    np.savez_compressed(os.path.join(HERE, "q15_mfcc.npz"), **out)

    rng = np.random.default_rng(20)
    x = rng.integers(-128, 128, (256, 403)).astype(np.int8)
    x[:64] //= 8
    x[64:80] = 0
    x[80:84] = 127
    x[84:88] = -128
    # MFCC-derived inputs: the golden frames above, tiled into 31-frame windows
    ni = np.clip(out["mfcc_simd"][:, :13], -128, 127).astype(np.int8)
    reps = (31 * 8 + ni.shape[0] - 1) // ni.shape[0]
    x[88:96] = np.tile(ni, (reps, 1))[: 31 * 8].reshape(8, 403)
    net = {"net_in": x}
    for name in ("local", "cmsis", "cmsis_simd"):
        L = ctypes.CDLL(os.path.join(REF, "libref_nnom_%s.so" % name))
        lg = np.zeros((x.shape[0], 10), np.int8); sm = np.zeros((x.shape[0], 10), np.int8)
        assert L.ref_nnom_run_batch(P(x), ctypes.c_long(x.shape[0]), P(lg), P(sm)) == 0
        net["logits_" + name] = lg
        net["softmax_" + name] = sm
    # adversarial softmax inputs: spreads beyond the 8-bit wrap of the SIMD branch
    adv = np.array([[127, -128, 0, 0, 0, 0, 0, 0, 0, 0], [127, -128, -128, -128, -128, -128, -128, -128, -128, -128],
                    [-128] * 10, [127] * 10, [0, 1, 2, 3, 4, 5, 6, 7, 8, 9], [100, -100, 50, -50, 25, -25, 12, -12, 6, -6],
                    [-128, -128, -128, -128, 127, 127, 127, 127, 0, 0], [5, 5, 5, 5, 5, 5, 5, 5, 5, 6]], dtype=np.int8)
    net["softmax_adv_in"] = adv
    for name in ("local", "cmsis", "cmsis_simd"):
        L = ctypes.CDLL(os.path.join(REF, "libref_nnom_%s.so" % name))
        o = np.zeros_like(adv)
        for i in range(adv.shape[0]):
            L.ref_softmax_q7(P(adv[i]), ctypes.c_int(10), P(o[i]))
        net["softmax_adv_" + name] = o
    np.savez_compressed(os.path.join(HERE, "nnom_q7.npz"), **net)
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
