"""Float CNN: the fp64 restatement (oracle/cube_float.py) on CPU, and K4 against it on the GPU.
Parity with keras / ST's runtime is UNPINNED (no stored tensors, closed runtime); what is checked here:
  * the blob round-trips the reference's weight array (sizes/offsets of kws.c:488-516),
  * the restatement classifies the reference's own recording of "edison" as class 0 "edison" with p > 0.9,
    as the int8 net does (a weak anchor for layout/order mistakes),
  * K4 == restatement within max-abs 1e-4 on softmax and 1e-3 relative on logits."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import cube_float as cf  # noqa: E402

BLOB = os.path.join(ROOT, "edison_b200", "data", "kws_cube_f32.bin")


def mfcc_like_inputs(oracle, n, seed):
    """float net inputs as the firmware forms them: q15 MFCC rows (first 13 coefficients) cast to float (app.c:679-683)"""
    from common import synth_frames
    pcm = synth_frames(n * 31, seed=seed)
    return oracle.mfcc(pcm)[:, :13].astype(np.float32).reshape(n, 403)


def test_blob_layout():
    L = cf.load(BLOB)
    assert [l["w"].shape for l in L] == [(16, 5, 5, 1), (32, 3, 3, 16), (64, 3, 3, 32), (32, 3, 3, 64), (10, 1, 1, 96)]
    assert sum(l["w"].size + l["b"].size for l in L) * 4 == 171944 - 0     # kws.c: AI_BUFFER 171944 bytes, no gaps but 5 pads
    assert os.path.getsize(BLOB) == 16 + 5 * 28 + 171944


def test_restatement_recognises_reference_recording(oracle):
    g = np.load(os.path.join(ROOT, "tests", "golden", "mfcc_float.npz"))
    pcm = np.zeros(32000, dtype=np.int16)                    # fileInference pads to 2 s with zeros (kws_on_mcu.py:287-288)
    pcm[:g["wav_pcm"].size] = g["wav_pcm"]
    mf = oracle.mfcc(pcm[:31 * 1024])
    _, sm = cf.forward(cf.load(BLOB), mf[:, :13].astype(np.float64).reshape(1, 403))
    assert sm.argmax() == 0 and sm[0, 0] > 0.9
    _, sm8 = oracle.forward(oracle.net_input(mf).reshape(1, 403))
    assert sm8.argmax() == 0


def test_restatement_fp32_vs_fp64(oracle):
    x = mfcc_like_inputs(oracle, 16, 3)
    L = cf.load(BLOB)
    _, s64 = cf.forward(L, x)
    _, s32 = cf.forward(L, x, dtype=np.float32)
    assert np.abs(s64 - s32).max() < 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 11, 12, 13, 500])
def test_gpu_float_cnn_vs_restatement(edison, oracle, n):
    x = mfcc_like_inputs(oracle, n, 40 + n)
    if n == 500:
        x[:50] = np.random.default_rng(1).integers(-300, 300, (50, 403)).astype(np.float32)
        x[50] = 0
    lg, sm = edison.infer_f32_batch(x)
    wl, ws = cf.forward(cf.load(BLOB), x)
    assert np.abs(sm - ws).max() <= 1e-4
    assert np.abs(lg - wl).max() <= 1e-3 * max(1.0, np.abs(wl).max())
    assert np.array_equal(sm.argmax(axis=1), ws.argmax(axis=1)) or np.sort(ws, axis=1)[:, -1].min() - np.sort(ws, axis=1)[:, -2].max() < 1e-3


@pytest.mark.gpu
def test_gpu_float_cnn_empty_and_bad_args(edison):
    lg, sm = edison.infer_f32_batch(np.zeros((0, 403), dtype=np.float32))
    assert lg.shape == (0, 10)
    assert edison.lib().edison_infer_f32_batch(None, 3, None, None) == -1
