"""-m "not gpu": the C-ABI library loads, exports every symbol include/edison_b200.h declares, fails loudly
without a GPU, and the host emulation of the K1 warp program agrees with the oracle."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

import common

ROOT = common.ROOT


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "edison_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    names = re.findall(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\([^;{}]*\)\s*;", txt)
    return sorted(set(n for n in names if n not in ("defined",)))


def test_library_exports_every_declared_symbol():
    import edison_b200
    L = edison_b200.lib()
    syms = declared_symbols()
    assert "audioCalcMFCCs" in syms and "edison_kws_batch" in syms and len(syms) >= 28
    for s in syms:
        assert hasattr(L, s), "missing export: " + s


def test_no_symbols_leak():
    out = subprocess.run(["nm", "-D", "--defined-only", os.path.join(ROOT, "edison_b200", "libedison_b200.so")],
                         capture_output=True, text=True).stdout
    names = [l.split()[-1] for l in out.splitlines() if l.strip()]
    allowed = set(declared_symbols()) | {"edison_internal_keyword", "edison_internal_keyword_count", "edison_internal_input_quant"}
    assert set(names) <= allowed, set(names) - allowed


def test_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import edison_b200
    with pytest.raises(edison_b200.EdisonError, match="no usable CUDA device|no CPU path"):
        edison_b200.init()
    with pytest.raises(edison_b200.EdisonError):
        edison_b200.mfcc_q15_batch(np.zeros((1, 1024), np.int16))


def test_product_does_not_link_oracle():
    out = subprocess.run(["ldd", os.path.join(ROOT, "edison_b200", "libedison_b200.so")], capture_output=True, text=True).stdout
    assert "oracle" not in out
    for dirpath, _, files in os.walk(os.path.join(ROOT, "edison_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".c", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "libedison_oracle" not in src and "oracle/" not in src.replace("the oracle", ""), f


@pytest.fixture(scope="module")
def emu():
    so = os.path.join(ROOT, "tests", "host_emu", "libemu_mfcc.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "-C", ROOT, "emu"])
    L = ctypes.CDLL(so)
    assert L.emu_init(os.path.join(ROOT, "edison_b200", "data", "mel_q15.bin").encode()) == 0
    return L


@pytest.mark.parametrize("force_sat", [-1, 1])
def test_host_emulation_of_k1_matches_oracle(emu, oracle, force_sat):
    """the CUDA source's index arithmetic (exchange layouts, twiddle tables, bit reversal, DCT batching),
    compiled for the host, against the oracle -- incl. a ragged group (n not a multiple of 8)"""
    pcm = np.concatenate([common.synth_frames(203, seed=31), common.lcg_probe_frames()])
    n = pcm.shape[0]
    mf = np.zeros((n, 32), np.int16)
    ni = np.zeros((n, 13), np.int8)
    emu.emu_mfcc_batch(common._p(pcm), ctypes.c_long(n), common._p(mf), common._p(ni), ctypes.c_int(force_sat))
    want = oracle.mfcc(pcm)
    assert (mf == want).all()
    assert (ni == oracle.net_input(want)).all()


def test_no_sat_path_is_safe_below_threshold(emu, oracle):
    """frames whose peak is <= 32000 must give the same bits with and without the saturating adds"""
    t = np.arange(1024)
    rows = [np.full(1024, 32000), np.full(1024, -32000), 32000 * np.where(t % 2 == 0, 1, -1)]
    for k in (1, 3, 64, 100, 255, 256, 300, 511):
        rows.append(32000 * np.sin(2 * np.pi * k / 1024 * t))
        rows.append(32000 * np.sign(np.sin(2 * np.pi * k / 1024 * t + 0.3)))
    rng = np.random.default_rng(4)
    rows += [rng.choice([-32000, 32000], 1024) for _ in range(20)]
    pcm = np.rint(np.stack(rows)).astype(np.int16)
    a = np.zeros((pcm.shape[0], 32), np.int16); b = a.copy()
    emu.emu_mfcc_batch(common._p(pcm), ctypes.c_long(pcm.shape[0]), common._p(a), None, ctypes.c_int(0))
    emu.emu_mfcc_batch(common._p(pcm), ctypes.c_long(pcm.shape[0]), common._p(b), None, ctypes.c_int(1))
    assert (a == b).all()
    assert (a == oracle.mfcc(pcm)).all()
