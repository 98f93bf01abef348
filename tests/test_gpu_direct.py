"""-m gpu: parity evidence that does not pass through the restatement oracle.

(a) the committed goldens (outputs of the reference's own C, tests/golden/make_golden.py) are fed to the CUDA path;
(b) when oracle/_ref/*.so travelled to the box (it is built from the unmodified reference sources by oracle/build_ref.sh), K1 and K2
    are compared with ref_mfcc_batch / ref_nnom_run_batch directly on fresh random inputs, and the restatement oracle is checked
    against the same outputs in the same test, so that this suite alone closes the chain GPU == oracle == reference;
(c) the checksum of the exact magnitudes of ALL 2^31 + 1 arguments, computed on the device, equals the one computed from the
    reference's arm_sqrt_q31 (tests/golden/make_mag_checksum.py)."""
import ctypes
import os

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu
GOLD = os.path.join(common.ROOT, "tests", "golden")
REF = os.path.join(common.ROOT, "oracle", "_ref")
has_ref = os.path.exists(os.path.join(REF, "libref_mfcc_simd.so")) and os.path.exists(os.path.join(REF, "libref_nnom_cmsis_simd.so"))
P = lambda a: a.ctypes.data_as(ctypes.c_void_p)

# sum over s in [0, 2^31] of (uint32)(int16)(arm_sqrt_q31(s) >> 16) * ((s & 1023) + 1) mod 2^64, from the reference's arm_sqrt_q31.c
MAG_SWEEP_CHECKSUM = 0x00556a2b2da19beb


def test_golden_q15_mfcc_on_gpu(edison):
    """probe frames, both reference WAVs, the hey_short single-frame fixture (audio/edison/mfcc/mfcc_on_mcu.py:321-323) and 96
    synthetic frames: the firmware build's audioCalcMFCCs outputs, and the net-input rows derived from them"""
    g = np.load(os.path.join(GOLD, "q15_mfcc.npz"))
    pcm, want = g["pcm"], g["mfcc_simd"]
    got = edison.mfcc_q15_batch(pcm)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert bad.size == 0, "golden frames %s differ" % bad[:8]
    # the same frames through the clip entry point (net-input rows = clip(mfcc[0:13], -128, 127), app.c:684-693)
    _, ni = edison.mfcc_q15_clips(pcm.reshape(-1), pcm.shape[0], 1, 1024, want_mfcc=False)
    assert (ni == np.clip(want[:, :13], -128, 127).astype(np.int8)).all()
    # the firmware symbol, frame by frame, on the WAV frames
    L = edison.capi.lib()
    res = ctypes.POINTER(ctypes.c_int16)()
    for i in range(3, 14):
        frame = pcm[i].copy()
        L.audioCalcMFCCs(P(frame), ctypes.byref(res))
        assert (np.ctypeslib.as_array(res, shape=(32,)) == want[i]).all(), i


def test_golden_nnom_q7_on_gpu(edison):
    """logits and softmax of the three reference NNoM builds (local backend, CMSIS-NN reference branch, CMSIS-NN SIMD branch =
    the firmware), incl. the adversarial softmax spreads that separate the SIMD branch from the reference branch"""
    g = np.load(os.path.join(GOLD, "nnom_q7.npz"))
    x = g["net_in"]
    try:
        for name, mode in (("cmsis", edison.SOFTMAX_CMSIS_REF), ("cmsis_simd", edison.SOFTMAX_CMSIS_SIMD), ("local", edison.SOFTMAX_NNOM_LOCAL)):
            edison.set_softmax_mode(mode)
            lg, sm = edison.infer_q7_batch(x)
            assert (lg == g["logits_" + name]).all(), name
            assert (sm == g["softmax_" + name]).all(), name
    finally:
        edison.set_softmax_mode(edison.SOFTMAX_CMSIS_SIMD)
    # default mode of the drop-in symbols = the firmware build (SIMD branch)
    out = np.zeros(10, np.int8)
    for i in (0, 70, 90, 255):
        assert edison.capi.lib().aiRunInference(P(x[i].copy()), P(out)) == 0
        assert (out == g["softmax_cmsis_simd"][i]).all(), i


@pytest.mark.skipif(not has_ref, reason="oracle/_ref not on this box (built from /root/reference by oracle/build_ref.sh)")
def test_k1_vs_reference_build_directly(edison, oracle):
    pcm = common.synth_frames(12000, seed=4242)
    L = ctypes.CDLL(os.path.join(REF, "libref_mfcc_simd.so"))
    want = np.zeros((pcm.shape[0], 32), np.int16)
    L.ref_mfcc_batch(P(pcm), ctypes.c_long(pcm.shape[0]), P(want))
    got = edison.mfcc_q15_batch(pcm)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert bad.size == 0, "frames %s differ from the reference build" % bad[:8]
    assert (oracle.mfcc(pcm[:3000]) == want[:3000]).all(), "the restatement oracle drifted from the reference build"


@pytest.mark.skipif(not has_ref, reason="oracle/_ref not on this box (built from /root/reference by oracle/build_ref.sh)")
def test_k2_vs_reference_builds_directly(edison, oracle):
    rng = np.random.default_rng(4243)
    n = 2400
    x = rng.integers(-128, 128, (n, 403)).astype(np.int8)
    x[:600] //= 8
    x[600:700] //= 32
    x[700:720] = 127
    x[720:740] = -128
    # MFCC-derived windows too
    ni = oracle.net_input(oracle.mfcc(common.synth_frames(31 * 40, seed=5))).reshape(40, 403)
    x[740:780] = ni
    try:
        for name, mode in (("cmsis", 0), ("cmsis_simd", 1), ("local", 2)):
            N = ctypes.CDLL(os.path.join(REF, "libref_nnom_%s.so" % name))
            wl = np.zeros((n, 10), np.int8); ws = np.zeros((n, 10), np.int8)
            assert N.ref_nnom_run_batch(P(x), ctypes.c_long(n), P(wl), P(ws)) == 0
            edison.set_softmax_mode(mode)
            lg, sm = edison.infer_q7_batch(x)
            assert (lg == wl).all(), name
            assert (sm == ws).all(), name
            ol, os_ = oracle.forward(x[:800], mode)
            assert (ol == wl[:800]).all() and (os_ == ws[:800]).all(), "the restatement oracle drifted from the reference build (%s)" % name
    finally:
        edison.set_softmax_mode(edison.SOFTMAX_CMSIS_SIMD)


def test_mag_sweep_checksum_vs_reference_sqrt(edison):
    """all 2^31 + 1 arguments of cmpl_mag_q15's arm_sqrt_q31: zero fast-path mismatches AND the exact path's checksum equals the
    reference's (arm_sqrt_q31.c:50-135 through oracle/_ref, constant made by tests/golden/make_mag_checksum.py)"""
    mism = ctypes.c_ulonglong(1)
    chk = ctypes.c_ulonglong(0)
    edison.capi.check(edison.capi.lib().edison_debug_mag_sweep(0, 0x80000001, ctypes.byref(mism), ctypes.byref(chk)))
    assert mism.value == 0
    assert chk.value == MAG_SWEEP_CHECKSUM, hex(chk.value)
