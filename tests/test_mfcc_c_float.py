"""firmware/src/audio/mfcc.c: the numpy restatement against the unmodified reference built for the host
(tests/golden/mfcc_c_float.npz from oracle/_ref/libref_mfcc_f32.so), and K3's mfcc.c mode against the same fixture.
Tolerance: q7 outputs equal up to ties of round(): <= 1 LSB, on frames none of whose 26 bands is (nearly) empty.  Frames of
constant / alternating full-scale samples are excluded: their upper bands hold only the float32 rounding noise of the
reference's own FFT, which no other implementation reproduces."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import mfcc_c_float as mc  # noqa: E402

TAGS = ["app", "dec2", "f1024", "f400"]


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "mfcc_c_float.npz"))


def ordinary_frames(pcm, n, flen, hop):
    ok = np.zeros(n, dtype=bool)
    for f in range(n):
        seg = pcm[f * hop:f * hop + flen].astype(np.int64)
        ok[f] = np.unique(np.abs(seg)).size > 2          # not constant, not +-alternating
    return ok


def check(got, want, ok):
    d = np.abs(got.astype(int) - want.astype(int))[ok]
    assert d.max(initial=0) <= 1, "max diff %d LSB" % d.max()
    assert (d > 0).mean() < 0.02, "%.3f of the outputs differ by 1 LSB" % (d > 0).mean()


@pytest.mark.parametrize("tag", TAGS)
def test_restatement_vs_reference_build(gold, tag):
    num, off, flen, dec, hop = (int(v) for v in gold["cfg_" + tag])
    m = mc.MfccC(num, off, flen, dec, 0.97)
    for src in ("wav", "syn"):
        pcm, want = gold[src + "_pcm"], gold["%s_%s" % (src, tag)]
        got, _ = m.batch(pcm, want.shape[0], hop)
        ok = ordinary_frames(pcm, want.shape[0], flen, hop)
        assert ok.sum() >= 0.8 * ok.size
        check(got, want, ok)
    # the reference WAV reproduces exactly
    assert np.array_equal(m.batch(gold["wav_pcm"], gold["wav_" + tag].shape[0], hop)[0], gold["wav_" + tag])


@pytest.mark.gpu
@pytest.mark.parametrize("tag", TAGS)
def test_gpu_mfcc_c_vs_reference_build(edison, gold, tag):
    num, off, flen, dec, hop = (int(v) for v in gold["cfg_" + tag])
    with edison.MfccC(num, off, flen, dec, 0.97) as m:
        assert m.n_out == num - off
        for src in ("wav", "syn"):
            pcm, want = gold[src + "_pcm"], gold["%s_%s" % (src, tag)]
            got = m.compute(pcm, want.shape[0], hop)
            check(got, want, ordinary_frames(pcm, want.shape[0], flen, hop))


@pytest.mark.gpu
def test_gpu_firmware_mfcc_symbols(edison, gold):
    import ctypes
    L = edison.lib()
    h = L.mfcc_create(13, 1, 512, 8, 0.97)      # app.c:540
    assert h
    pcm = np.ascontiguousarray(gold["wav_pcm"][256 * 7:256 * 7 + 512])
    out = np.zeros(12, dtype=np.int8)
    L.mfcc_compute(h, pcm.ctypes.data_as(ctypes.c_void_p), out.ctypes.data_as(ctypes.c_void_p))
    L.mfcc_delete(h)
    assert np.abs(out.astype(int) - gold["wav_app"][7].astype(int)).max() <= 1
    bad = ctypes.c_void_p()
    assert L.edison_mfccf_create(13, 1, 4096, 8, 0.97, ctypes.byref(bad)) == -1


@pytest.mark.gpu
def test_gpu_mfcc_c_odd_hop_and_length(edison):
    # odd hop / odd frame length take the unaligned loader; compared with the restatement
    from common import synth_frames
    pcm = synth_frames(6, seed=3).reshape(-1)
    for flen, hop in ((511, 255), (512, 333), (300, 1)):
        n = min(40, 1 + (pcm.size - flen) // hop)
        want, _ = mc.MfccC(13, 0, flen, 2, 0.97).batch(pcm, n, hop)
        with edison.MfccC(13, 0, flen, 2, 0.97) as m:
            got = m.compute(pcm, n, hop)
        check(got, want, ordinary_frames(pcm, n, flen, hop))
