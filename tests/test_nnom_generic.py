"""Generic NNoM executor (include/edison_b200.h section 2d).
Pin: the upstream NNoM keyword_spotting example model run through the reference's unmodified NNoM runtime built for the
host (tests/golden/nnom_example.npz, made by tests/golden/make_golden_nnom_example.py).  CPU: the oracle's generic forward
against that fixture.  GPU: edison_model_* against the fixture, against the oracle on random models of the same layer family,
and against K2 on the reference's own model."""
import ctypes
import os
import struct

import numpy as np
import pytest

import common

ROOT = common.ROOT
GOLD = os.path.join(ROOT, "tests", "golden")
EXAMPLE = os.path.join(GOLD, "nnom_kws_example.bin")


def oracle_generic(oracle, blob, x, mode=0):
    """run the oracle's generic forward on `blob`, then put the reference model back (the oracle holds one model)"""
    L = oracle.L
    try:
        assert L.eo_nnom_load(os.fsencode(blob)) == 0
        ni, no = L.eo_nnom_in_size(), L.eo_nnom_out_size()
        x = np.ascontiguousarray(x, dtype=np.int8).reshape(-1, ni)
        lg = np.zeros((x.shape[0], no), dtype=np.int8)
        sm = np.zeros((x.shape[0], no), dtype=np.int8)
        assert L.eo_nnom_forward_generic(common._p(x), ctypes.c_size_t(x.shape[0]), common._p(lg), common._p(sm), ctypes.c_int(mode)) == 0
        return lg, sm
    finally:
        assert L.eo_nnom_load(common.BLOB.encode()) == 0


def random_blob(path, rng, in_shape, convs, dense_out):
    """convs: list of (cout, kh, kw, relu, pool).  Weights small enough that activations stay lively."""
    h, w, c = in_shape
    out = bytearray(b"EDNNQ7\0\0" + struct.pack("<II", 1, len(convs) + 1))
    for cout, kh, kw, relu, pool in convs:
        wt = rng.integers(-40, 40, cout * kh * kw * c).astype(np.int8)
        b = rng.integers(-60, 60, cout).astype(np.int8)
        out += struct.pack("<9I", 1, cout, kh, kw, c, int(rng.integers(5, 9)), int(rng.integers(0, 5)), relu, pool) + wt.tobytes() + b.tobytes()
        out += b"\0" * (-len(out) % 4)
        h, w, c = h - kh + 1, w - kw + 1, cout
        if pool:
            h = (h - pool) // pool + 1
    n_in = h * w * c
    wt = rng.integers(-30, 30, dense_out * n_in).astype(np.int8)
    b = rng.integers(-60, 60, dense_out).astype(np.int8)
    out += struct.pack("<9I", 2, dense_out, 1, 1, n_in, int(rng.integers(6, 10)), int(rng.integers(0, 4)), 0, 0) + wt.tobytes() + b.tobytes()
    out += b"\0" * (-len(out) % 4)
    out += struct.pack("<iiiIII", 1, -128, 127, *in_shape)
    open(path, "wb").write(out)


def test_oracle_generic_vs_reference_runtime(oracle):
    g = np.load(os.path.join(GOLD, "nnom_example.npz"))
    lg, sm = oracle_generic(oracle, EXAMPLE, g["x"], mode=2)     # NNoM local backend -> local_softmax_q7
    assert np.array_equal(lg, g["logits"])
    assert np.array_equal(sm, g["softmax_local"])
    assert len(set(map(bytes, g["logits"]))) > 40                # the fixture is not degenerate


@pytest.mark.gpu
def test_gpu_generic_vs_reference_runtime(edison):
    g = np.load(os.path.join(GOLD, "nnom_example.npz"))
    edison.set_softmax_mode(edison.SOFTMAX_NNOM_LOCAL)
    try:
        with edison.Model(EXAMPLE) as m:
            assert m.in_shape == (63, 12, 1) and m.n_out == 35
            lg, sm = m.infer(g["x"])
    finally:
        edison.set_softmax_mode(edison.SOFTMAX_CMSIS_SIMD)
    assert np.array_equal(lg, g["logits"])
    assert np.array_equal(sm, g["softmax_local"])


@pytest.mark.gpu
@pytest.mark.parametrize("case", range(4))
def test_gpu_generic_random_models_vs_oracle(edison, oracle, tmp_path, case):
    rng = np.random.default_rng(100 + case)
    spec = [((20, 9, 1), [(8, 3, 3, 1, 2), (12, 3, 3, 1, 0)], 7),
            ((17, 11, 3), [(5, 2, 4, 1, 3), (16, 3, 1, 0, 0), (8, 1, 3, 1, 2)], 12),
            ((31, 13, 1), [(16, 5, 5, 1, 2), (32, 3, 3, 1, 2), (64, 3, 3, 1, 0), (32, 3, 3, 1, 0)], 10),
            ((12, 12, 4), [(8, 3, 3, 1, 0)], 33)][case]
    blob = str(tmp_path / "m.bin")
    random_blob(blob, rng, *spec)
    n = 37
    x = rng.integers(-128, 128, (n, int(np.prod(spec[0])))).astype(np.int8)
    for mode in (0, 1, 2):
        want_lg, want_sm = oracle_generic(oracle, blob, x, mode)
        edison.set_softmax_mode(mode)
        try:
            with edison.Model(blob) as m:
                lg, sm = m.infer(x)
        finally:
            edison.set_softmax_mode(edison.SOFTMAX_CMSIS_SIMD)
        assert np.array_equal(lg, want_lg), "logits, mode %d" % mode
        assert np.array_equal(sm, want_sm), "softmax, mode %d" % mode
    assert np.unique(want_lg).size > 8


@pytest.mark.gpu
def test_gpu_generic_equals_k2_on_reference_model(edison, oracle):
    x = oracle.net_input(oracle.mfcc(common.synth_frames(31 * 40, seed=8))).reshape(40, 403)
    with edison.Model(common.BLOB) as m:
        lg, sm = m.infer(x)
    lg2, sm2 = edison.infer_q7_batch(x)
    assert np.array_equal(lg, lg2) and np.array_equal(sm, sm2)


@pytest.mark.gpu
def test_gpu_generic_rejects_bad_blobs(edison, tmp_path):
    L = edison.lib()
    h = ctypes.c_void_p()
    assert L.edison_model_load(b"/nonexistent.bin", ctypes.byref(h)) == -4
    p = tmp_path / "bad.bin"
    p.write_bytes(b"EDNNQ7\0\0" + struct.pack("<II", 1, 1) + struct.pack("<9I", 2, 10, 1, 1, 999, 8, 0, 0, 0) + bytes(9990 + 10 + 24 + 8))
    assert L.edison_model_load(os.fsencode(str(p)), ctypes.byref(h)) == -6      # dense input does not match the 0x0x0 input
    # header whose weight count wraps size_t: cout * kh * kw * cin = 2^32 * 2^32; refused as malformed data, nothing dereferenced
    p2 = tmp_path / "hostile.bin"
    p2.write_bytes(b"EDNNQ7\0\0" + struct.pack("<II", 1, 1) + struct.pack("<9I", 1, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 8, 0, 0, 0) + bytes(64))
    assert L.edison_model_load(os.fsencode(str(p2)), ctypes.byref(h)) == -4
    p3 = tmp_path / "hostile2.bin"      # 65536^4 = 2^64 wraps to 0
    p3.write_bytes(b"EDNNQ7\0\0" + struct.pack("<II", 1, 1) + struct.pack("<9I", 1, 65536, 65536, 65536, 65536, 8, 0, 0, 0) + bytes(64))
    assert L.edison_model_load(os.fsencode(str(p3)), ctypes.byref(h)) == -4
