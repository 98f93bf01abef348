"""-m gpu: the CUDA path through the C ABI against the CPU oracle, bit-exact (q15 / int8 / index work)."""
import ctypes

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


def test_probe_vectors(edison, oracle):
    """the firmware-true (SIMD build) outputs of the three LCG probe frames."""
    out = edison.mfcc_q15_batch(common.lcg_probe_frames())
    expect0 = [5176, -2317, 616, -327, 119, 171, -9, -103, 159, -25, -10, 7, 42, -17, 7, 7, 38]
    expect1 = [637, -297, 81, -49, 28, -20, 26, -15, 6, 4, -9, -6, -3, 15, -4, -9, 4]
    expect2 = [74, -34, 7, -9, 5, -4, -1, 1, 0, -1, 0, -2, 3, -1, 0, -3, 4]
    assert out[0, :17].tolist() == expect0
    assert out[1, :17].tolist() == expect1
    assert out[2, :17].tolist() == expect2
    for k in range(17, 32):
        assert (out[:, k] == out[:, 32 - k]).all()


@pytest.mark.parametrize("n", [1, 7, 8, 9, 257, 6000])
def test_mfcc_q15_vs_oracle(edison, oracle, n):
    pcm = common.synth_frames(n, seed=20 + n)
    got = edison.mfcc_q15_batch(pcm)
    want = oracle.mfcc(pcm)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert bad.size == 0, "frames %s differ, first: got %s want %s" % (bad[:8], got[bad[0]], want[bad[0]])


def test_mfcc_q15_launch_geometries(edison, oracle):
    """K1 starts whole 4-warp barrier groups, no more than the work needs, and warps without work compute on stale registers:
    every warp count (groups of 8 frames: 1..21 groups in one CTA, then two CTAs) and ragged last groups must stay exact
    and must not hang on a barrier."""
    pcm = common.synth_frames(8 * 45 + 5, seed=77)
    want = oracle.mfcc(pcm)
    for n in list(range(1, 34)) + [8 * k + r for k in (4, 8, 12, 16, 19, 20, 21, 40, 44) for r in (0, 1, 7)] + [pcm.shape[0]]:
        got = edison.mfcc_q15_batch(pcm[:n])
        assert (got == want[:n]).all(), n


def test_mfcc_q15_repeatable(edison, oracle):
    """warp groups drift against each other inside a CTA (named barriers per group): repeated launches on the same input must
    give the same bits every time, and the oracle's"""
    pcm = common.synth_frames(20000, seed=91)
    first = edison.mfcc_q15_batch(pcm)
    for _ in range(4):
        assert (edison.mfcc_q15_batch(pcm) == first).all()
    pick = np.arange(0, 20000, 37)
    assert (first[pick] == oracle.mfcc(pcm[pick])).all()


def test_mfcc_empty(edison):
    assert edison.mfcc_q15_batch(np.zeros((0, 1024), np.int16)).shape == (0, 32)


def test_net_input_rows_and_ragged_clips(edison, oracle):
    """1 s clips (16,000 samples -> 15 frames, stride 16,000: config #2's framing) with odd-frame tail."""
    n_clips, fpc, stride = 37, 15, 16000
    rng = np.random.default_rng(5)
    pcm = (rng.normal(0, 4000, n_clips * stride)).clip(-32768, 32767).astype(np.int16)
    mf, ni = edison.mfcc_q15_clips(pcm, n_clips, fpc, stride)
    frames = np.stack([pcm[c * stride + f * 1024: c * stride + (f + 1) * 1024] for c in range(n_clips) for f in range(fpc)])
    want = oracle.mfcc(frames)
    assert (mf == want).all()
    assert (ni == oracle.net_input(want)).all()


def test_odd_sample_stride(edison, oracle):
    """clip stride that is not a multiple of two samples exercises the 16-bit load path"""
    n_clips, fpc, stride = 9, 3, 3 * 1024 + 1
    rng = np.random.default_rng(6)
    pcm = rng.integers(-3000, 3000, n_clips * stride + 8).astype(np.int16)
    mf, _ = edison.mfcc_q15_clips(pcm, n_clips, fpc, stride, want_net_in=False)
    frames = np.stack([pcm[c * stride + f * 1024: c * stride + (f + 1) * 1024] for c in range(n_clips) for f in range(fpc)])
    assert (mf == oracle.mfcc(frames)).all()


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_infer_q7_vs_oracle(edison, oracle, mode):
    rng = np.random.default_rng(11 + mode)
    x = rng.integers(-128, 128, (1003, 403)).astype(np.int8)
    x[:200] //= 8
    x[200:300] = 0
    x[300:310] = 127
    x[310:320] = -128
    edison.set_softmax_mode(mode)
    try:
        lg, sm = edison.infer_q7_batch(x)
    finally:
        edison.set_softmax_mode(edison.SOFTMAX_CMSIS_SIMD)
    wl, ws = oracle.forward(x, mode)
    assert (lg == wl).all()
    assert (sm == ws).all()


def test_kws_end_to_end(edison, oracle):
    n = 64
    pcm = common.synth_frames(n * 31, seed=3).reshape(-1)
    lg, sm, ni = edison.kws_batch(pcm, n, want_net_in=True)
    wl, ws, wn = oracle.kws(pcm, n)
    assert (ni == wn).all()
    assert (lg == wl).all()
    assert (sm == ws).all()


def test_sqrt_q31_device_vs_oracle(edison, oracle):
    """arm_sqrt_q31 on the device (exact Newton path) against the oracle, incl. edge values"""
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.integers(0, 2**31, 200000), rng.integers(0, 70000, 20000),
                        np.array([0, 1, 2, 3, 4, 2**31 - 1, 2**30, 2**29, -1, -2**31, 18, 72, 98, 128])]).astype(np.int32)
    out = np.empty_like(x)
    edison.capi.check(edison.capi.lib().edison_debug_sqrt_q31(x.ctypes.data_as(ctypes.c_void_p), x.size,
                                                         out.ctypes.data_as(ctypes.c_void_p)))
    want = np.array([oracle.L.eo_sqrt_q31(int(v)) for v in x], dtype=np.int32)
    assert (out == want).all()


def test_mag_fast_path_exhaustive(edison):
    """the fast magnitude path equals the exact Newton path for EVERY 32-bit sum re^2+im^2 can take"""
    mism = ctypes.c_ulonglong(0)
    chk = ctypes.c_ulonglong(0)
    edison.capi.check(edison.capi.lib().edison_debug_mag_sweep(0, 0x80000001, ctypes.byref(mism), ctypes.byref(chk)))
    assert mism.value == 0


def test_firmware_entry_points(edison, oracle):
    L = edison.capi.lib()
    frame = common.synth_frames(1, seed=9)[0].copy()
    res = ctypes.POINTER(ctypes.c_int16)()
    L.audioInit()
    L.audioCalcMFCCs(frame.ctypes.data_as(ctypes.c_void_p), ctypes.byref(res))
    got = np.ctypeslib.as_array(res, shape=(32,)).copy()
    assert (got == oracle.mfcc(frame)[0]).all()
    x = np.random.default_rng(2).integers(-128, 128, 403).astype(np.int8)
    out = np.zeros(10, np.int8)
    assert L.aiRunInference(x.ctypes.data_as(ctypes.c_void_p), out.ctypes.data_as(ctypes.c_void_p)) == 0
    assert (out == oracle.forward(x)[1][0]).all()
    assert L.aiGetKeywordCount() == 10
    assert L.aiGetKeywordFromIndex(0) == b"edison"
    xs, ys = ctypes.c_uint16(0), ctypes.c_uint16(0)
    L.aiGetInputShape(ctypes.byref(xs), ctypes.byref(ys))
    assert (xs.value, ys.value) == (13, 31)
