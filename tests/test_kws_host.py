"""Host helpers of edison_b200.kws (the fileInference / frameInference clip preparation of kws_on_mcu.py:273-345)."""
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _kws():
    import importlib
    return importlib.import_module("edison_b200.kws")


def test_fit_clip_cut_zero_edge():
    k = _kws()
    assert (k.SAMPLE_LEN, k.N_FRAMES) == (32000, 31)
    x = np.arange(1, 101, dtype=np.int16)
    z = k.fit_clip(x, "zero")
    e = k.fit_clip(x, "edge")
    assert z.shape == e.shape == (32000,) and z.dtype == e.dtype == np.int16
    assert np.array_equal(z[:100], x) and not z[100:].any()
    assert np.array_equal(e[:100], x) and (e[100:] == 100).all()
    long = np.arange(40000, dtype=np.int32).astype(np.int16)
    assert np.array_equal(k.fit_clip(long), long[:32000])
    assert np.array_equal(k.fit_clip(np.zeros(0, np.int16)), np.zeros(32000, np.int16))
    with pytest.raises(ValueError):
        k.fit_clip(np.zeros(0, np.int16), "edge")
    with pytest.raises(ValueError):
        k.fit_clip(x, "wrap")


def test_to_int16_float_scaling_truncates():
    k = _kws()
    f = np.array([0.0, 0.5, -0.5, 0.99999, -1.0, 1e-5], dtype=np.float32)
    want = ((2 ** 15 - 1) * f).astype("int16")              # the reference's expression, kws_on_mcu.py:197-198
    assert np.array_equal(k.to_int16(f), want) and want[1] == 16383 and want[2] == -16383
    with pytest.raises(TypeError):
        k.to_int16(np.zeros(4, dtype=np.float64))


@pytest.mark.gpu
def test_file_and_frame_inference_match_oracle(edison, oracle):
    k = _kws()
    g = np.load(os.path.join(ROOT, "tests", "golden", "mfcc_float.npz"))
    wav = g["wav_pcm"]                                       # the reference's own recording of "edison"
    for pad, fn in (("zero", k.file_inference), ("edge", k.frame_inference)):
        ni, lg, sm, cls = fn(wav)
        clip = k.fit_clip(wav, pad)
        mf = oracle.mfcc(clip[:31 * 1024])
        want_in = oracle.net_input(mf)
        want_lg, want_sm = oracle.forward(want_in.reshape(1, 403))
        assert np.array_equal(ni.reshape(-1), want_in.reshape(-1))
        assert np.array_equal(lg, want_lg[0]) and np.array_equal(sm, want_sm[0])
        assert cls == 0
    # a batch of differently padded clips in one call
    clips = np.stack([k.fit_clip(wav, "zero"), k.fit_clip(wav[:9000], "edge"), k.fit_clip(wav[3000:], "zero")])
    ni, lg, sm = k.infer_clips(clips)
    for i in range(3):
        w_in = oracle.net_input(oracle.mfcc(clips[i, :31 * 1024]))
        w_lg, w_sm = oracle.forward(w_in.reshape(1, 403))
        assert np.array_equal(ni[i].reshape(-1), w_in.reshape(-1)) and np.array_equal(lg[i], w_lg[0]) and np.array_equal(sm[i], w_sm[0])
