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
