"""The K2 tcgen05 plan (edison_b200/csrc/k2_umma_plan.h: shared-memory layout, MMA list, unit order, epilogues, weight image) run by a
host emulation (tests/host_emu/emu_k2.cpp) must give the oracle's bits.  This checks the index arithmetic of the CUDA
kernel's plan without a GPU; the -m gpu parity tests run the kernel itself."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import common

ROOT = common.ROOT


@pytest.fixture(scope="module")
def emu():
    so = os.path.join(ROOT, "tests", "host_emu", "libemu_k2.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "-C", ROOT, "emu"])
    L = ctypes.CDLL(so)
    L.emu_k2_image_bytes.restype = ctypes.c_uint
    L.emu_k2_smem_bytes.restype = ctypes.c_uint
    return L


def run(emu, x, mode, eager):
    n = x.shape[0]
    lg = np.zeros((n, 10), np.int8)
    sm = np.zeros((n, 10), np.int8)
    hist = np.zeros(10, np.int64)
    assert emu.emu_k2_run(common.BLOB.encode(), common._p(x), ctypes.c_longlong(n), common._p(lg), common._p(sm), mode,
                          common._p(hist), eager) == 0
    return lg, sm, hist


def test_plan_fits_a_cta(emu):
    assert emu.emu_k2_smem_bytes() <= 227 * 1024
    assert emu.emu_k2_image_bytes() % 16 == 0


@pytest.mark.parametrize("n", [1, 31, 32, 33, 97])
@pytest.mark.parametrize("eager", [0, 1])
def test_plan_matches_oracle(emu, oracle, n, eager):
    rng = np.random.default_rng(7 * n + eager)
    x = rng.integers(-128, 128, (n, 403), dtype=np.int8)
    x[1::3] //= 8           # quiet inputs: small accumulators, rounding and ReLU edges
    x[2::5] = 127           # saturating inputs
    for mode in (0, 1, 2):
        want_lg, want_sm = oracle.forward(x, mode)
        lg, sm, hist = run(emu, x, mode, eager)
        assert np.array_equal(lg, want_lg), "logits"
        assert np.array_equal(sm, want_sm), "softmax mode %d" % mode
        assert np.array_equal(hist, np.bincount(want_lg.argmax(axis=1), minlength=10)), "class histogram (first maximum)"


def test_plan_on_real_net_inputs(emu, oracle):
    """net inputs as K1 produces them (MFCC rows of synthetic audio), incl. all-zero and saturated clips"""
    pcm = common.synth_frames(31 * 40, seed=5)
    ni = oracle.net_input(oracle.mfcc(pcm)).reshape(40, 403)
    want_lg, want_sm = oracle.forward(ni, 1)
    lg, sm, _ = run(emu, ni, 1, 1)
    assert np.array_equal(lg, want_lg) and np.array_equal(sm, want_sm)


@pytest.mark.parametrize("eager", [0, 1])
def test_incremental_schedule_matches_full_window(emu, oracle, eager):
    """the incremental mode's own schedule (kSchedI: one pipeline stage per iteration, static slot map, row cache with four phase sets of
    three slots) emulated hop by hop on sliding windows: from the first inferring hop on, its softmax must equal the oracle's on the whole
    window.  Three batches on one CTA (the last one ragged), so that five batches' stages overlap; both execution orders."""
    n, hops = 70, 46
    rng = np.random.default_rng(31)
    rows = rng.integers(-128, 128, (hops + 31, n, 13)).astype(np.int8)      # row t of every stream; the window at hop h is rows h .. h + 30
    rows[:, 0, :] = 0
    emu.emu_k2_cache_batch_bytes.restype = ctypes.c_uint
    nb = (n + 31) // 32
    cache = np.zeros(nb * emu.emu_k2_cache_batch_bytes(), np.uint8)
    checked = 0
    for h in range(1, hops + 1):
        win = np.ascontiguousarray(rows[h:h + 31].transpose(1, 0, 2).reshape(n, 403))
        use = (h >> 2) % 3
        sm = np.full((n, 10), 99, np.int8)
        assert emu.emu_k2_incr_hop(common.BLOB.encode(), common._p(win), ctypes.c_longlong(n), common._p(cache), h & 3, (use + 1) % 3, (use + 2) % 3, use,
                                   common._p(sm), 1, eager) == 0
        if h >= 20:      # every cached row in use was computed from real rows by now (8 hops of history; stream.cu infers from hop 32 on)
            _, want = oracle.forward(win, 1)
            assert np.array_equal(sm, want), "hop %d" % h
            checked += 1
    assert checked == hops - 19
