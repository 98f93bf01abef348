"""The numpy restatement of the reference's float MFCCs (oracle/mfcc_float.py) against outputs of the reference module
itself (tests/golden/mfcc_float.npz, made by tests/golden/make_golden_float.py)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import mfcc_float as mf  # noqa: E402


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "mfcc_float.npz"))


def test_survey_probe_values(gold):
    # mfcc[0][:13] and mfcc[3][:13] of audio/data/edison_16k_16b.wav, mfcc_mcu[3][:13]
    want0 = [65.5856, -0.5113, 0.7374, -0.0945, 0.1716, -0.0291, -0.0009, 0.0917, 0.3237, 0.1907, 0.1312, 0.4222, 0.3843]
    want3 = [97.529, 2.4794, 0.5165, 2.815, 0.4109, -0.5967, -0.3244, 0.0177, 0.3177, 0.1132, 0.0123, -0.2744, 0.0618]
    mcu3 = [184.28, 65.52, 34.81, 66.06, 29.01, 0.37, 6.62, 7.77, 2.55, 3.58, 7.62, -0.48, 2.68]
    assert gold["wav_host_mfcc"].shape == (10, 32)
    assert np.abs(gold["wav_host_mfcc"][0][:13] - want0).max() < 6e-5
    assert np.abs(gold["wav_host_mfcc"][3][:13] - want3).max() < 6e-4
    assert np.abs(gold["wav_mcu_mfcc"][3][:13] - mcu3).max() < 6e-3


def test_mel_matrices(gold):
    for n in (512, 513):
        w = mf.mel_weight_matrix(n)
        assert np.array_equal(w, gold["mel_w%d" % n])
    w = gold["mel_w513"]
    assert (w > 0).sum(axis=0).max() <= 72 and w[0].sum() == 0


@pytest.mark.parametrize("src", ["wav", "syn"])
def test_restatement_equals_reference(gold, src):
    pcm = gold[src + "_pcm"]
    assert np.array_equal(mf.mfcc_host(pcm), gold[src + "_host_mfcc"])
    assert np.array_equal(mf.mfcc_mcu(pcm), gold[src + "_mcu_mfcc"])
    assert np.array_equal(mf.mfcc_mcu(pcm, use_log=True), gold[src + "_mculog_mfcc"])


def test_framing():
    assert mf.n_frames(11243) == 10 and mf.n_frames(16000) == 15 and mf.n_frames(31744) == 31 and mf.n_frames(1024) == 1
