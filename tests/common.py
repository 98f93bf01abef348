"""Shared test helpers: ctypes wrapper of the CPU oracle and the seeded synthetic inputs of SURVEY 8(d).
TEST INFRASTRUCTURE -- nothing here is imported by the product package."""
import ctypes
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BLOB = os.path.join(ROOT, "edison_b200", "data", "kws_nnom_q7.bin")
FRAME = 1024


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


class Oracle:
    """oracle/libedison_oracle.so (CPU restatement, pinned against oracle/_ref in tests/test_oracle_cpu.py).
    Default softmax mode = 1, the firmware build's SIMD branch of arm_softmax_q7, like edison_init(NULL)."""

    def __init__(self):
        self.L = ctypes.CDLL(os.path.join(ROOT, "oracle", "libedison_oracle.so"))
        self.L.eo_init()
        self.L.eo_sqrt_q31.restype = ctypes.c_int32
        self.L.eo_sqrt_q31.argtypes = [ctypes.c_int32]
        assert self.L.eo_nnom_load(BLOB.encode()) == 0

    def mfcc(self, pcm):
        pcm = np.ascontiguousarray(pcm, dtype=np.int16).reshape(-1, FRAME)
        out = np.zeros((pcm.shape[0], 32), dtype=np.int16)
        self.L.eo_mfcc_batch(_p(pcm), ctypes.c_size_t(pcm.shape[0]), _p(out))
        return out

    def net_input(self, mfcc):
        mfcc = np.ascontiguousarray(mfcc, dtype=np.int16).reshape(-1, 32)
        out = np.zeros((mfcc.shape[0], 13), dtype=np.int8)
        for i in range(mfcc.shape[0]):
            self.L.eo_net_input_row(_p(mfcc[i]), _p(out[i]))
        return out

    def forward(self, net_in, mode=1):
        net_in = np.ascontiguousarray(net_in, dtype=np.int8).reshape(-1, 403)
        n = net_in.shape[0]
        lg = np.zeros((n, 10), dtype=np.int8)
        sm = np.zeros((n, 10), dtype=np.int8)
        assert self.L.eo_nnom_forward_batch(_p(net_in), ctypes.c_size_t(n), _p(lg), _p(sm), ctypes.c_int(mode)) == 0
        return lg, sm

    def kws(self, pcm, n_clips, stride=31 * FRAME, mode=1):
        pcm = np.ascontiguousarray(pcm, dtype=np.int16).reshape(-1)
        lg = np.zeros((n_clips, 10), dtype=np.int8)
        sm = np.zeros((n_clips, 10), dtype=np.int8)
        ni = np.zeros((n_clips, 403), dtype=np.int8)
        assert self.L.eo_kws_batch(_p(pcm), ctypes.c_size_t(n_clips), ctypes.c_size_t(stride), _p(lg), _p(sm), _p(ni),
                                   ctypes.c_int(mode)) == 0
        return lg, sm, ni

    def softmax(self, x, mode):
        x = np.ascontiguousarray(x, dtype=np.int8)
        out = np.zeros_like(x)
        self.L.eo_softmax_q7(_p(x), ctypes.c_int(x.size), _p(out), ctypes.c_int(mode))
        return out


def lcg_probe_frames():
    """The three probe frames of SURVEY 8(c) (LCG, amplitude /1, /8, /64)."""
    s = 12345
    out = np.zeros((3, FRAME), dtype=np.int16)
    for t, div in enumerate((1, 8, 64)):
        for i in range(FRAME):
            s = (s * 1664525 + 1013904223) & 0xFFFFFFFF
            v = (s >> 16) - 32768
            out[t, i] = int(v / div)  # C division truncates toward zero
    return out


def synth_frames(n, seed=20):
    """n frames cycling through the distributions D1..D4, D6 of SURVEY 8(d); deterministic."""
    rng = np.random.default_rng(seed)
    t = np.arange(FRAME)
    out = np.zeros((n, FRAME), dtype=np.float64)
    for i in range(n):
        k = i % 6
        if k == 0:  # D1 speech-like: 4 sinusoids + noise
            x = rng.normal(0, 200, FRAME)
            for _ in range(4):
                x += rng.uniform(500, 8000) * np.sin(2 * np.pi * rng.uniform(100, 4000) / 16000 * t + rng.uniform(0, 6.28))
            out[i] = x
        elif k == 1:  # D2 full-scale uniform
            out[i] = rng.integers(-32768, 32768, FRAME)
        elif k == 2:  # D3 low level
            out[i] = rng.integers(-64, 65, FRAME)
        elif k == 3:  # D4 constants / Nyquist / impulses / full-scale tones
            j = (i // 6) % 8
            if j == 0: out[i] = 32767
            elif j == 1: out[i] = -32768
            elif j == 2: out[i, ::2], out[i, 1::2] = 32767, -32768
            elif j == 3: out[i, rng.integers(0, FRAME)] = 32767
            elif j == 4: out[i] = 0
            elif j == 5: out[i] = 32767 * np.sign(np.sin(2 * np.pi * rng.integers(1, 512) / FRAME * t + 0.1))
            elif j == 6: out[i] = 32767 * np.sin(2 * np.pi * rng.integers(1, 512) / FRAME * t)
            else: out[i] = rng.choice([-32768, 32767], FRAME)
        elif k == 4:  # mid-level uniform (never saturates)
            out[i] = rng.integers(-20000, 20001, FRAME)
        else:  # gaussian, occasionally clipping
            out[i] = rng.normal(0, rng.choice([300, 3000, 15000]), FRAME)
    return np.clip(np.rint(out), -32768, 32767).astype(np.int16)
