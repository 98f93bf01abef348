"""CPU baseline of the KWS hot path on the host cores (TEST / MEASUREMENT INFRASTRUCTURE).

Times the reference's own C (oracle/_ref: firmware/src/audioprocessing.c at -O0 + CMSIS-DSP SIMD
branches on host-emulated M4 intrinsics, NNoM + CMSIS-NN SIMD branches -- the firmware-true build,
SURVEY 8(c)) when oracle/_ref exists, else the oracle port (oracle/libedison_oracle.so).  The
reference code is not re-entrant (static buffers), so it is parallelised with one PROCESS per core
over disjoint clip ranges.  Only bench.py's cpu_baseline / --impl reference legs call this.
"""
import ctypes
import multiprocessing as mp
import os
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_MFCC = os.path.join(HERE, "_ref", "libref_mfcc_simd.so")
REF_NNOM = os.path.join(HERE, "_ref", "libref_nnom_cmsis_simd.so")
PORT = os.path.join(HERE, "libedison_oracle.so")
BLOB = os.path.join(os.path.dirname(HERE), "edison_b200", "data", "kws_nnom_q7.bin")


def kind():
    return "reference" if os.path.exists(REF_MFCC) and os.path.exists(REF_NNOM) else "port"


def _synth(n_clips, seed):
    """speech-like clips (sinusoids + noise), the D1 distribution of SURVEY 8(d)"""
    rng = np.random.default_rng(seed)
    n = n_clips * 31 * 1024
    t = np.arange(n, dtype=np.float64)
    x = rng.normal(0, 200, n)
    for _ in range(4):
        x += rng.uniform(500, 8000) * np.sin(2 * np.pi * rng.uniform(100, 4000) / 16000 * t + rng.uniform(0, 6.28))
    return np.clip(np.rint(x), -32768, 32767).astype(np.int16)


def _worker(args):
    n_clips, seed, mfcc_only, reps, which = args
    pcm = _synth(n_clips, seed)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    nf = n_clips * 31
    mf = np.zeros((nf, 32), np.int16)
    lg = np.zeros((n_clips, 10), np.int8)
    sm = np.zeros((n_clips, 10), np.int8)
    if which == "reference":
        A, N = ctypes.CDLL(REF_MFCC), ctypes.CDLL(REF_NNOM)
        N.ref_nnom_init()

        def step():
            A.ref_mfcc_batch(p(pcm), ctypes.c_long(nf), p(mf))
            if not mfcc_only:
                ni = np.clip(mf[:, :13], -128, 127).astype(np.int8).reshape(n_clips, 403)  # mfccToNetInput, scale 1
                N.ref_nnom_run_batch(p(ni), ctypes.c_long(n_clips), p(lg), p(sm))
    else:
        O = ctypes.CDLL(PORT)
        O.eo_init()
        O.eo_nnom_load(BLOB.encode())

        def step():
            if mfcc_only:
                O.eo_mfcc_batch(p(pcm), ctypes.c_size_t(nf), p(mf))
            else:
                O.eo_kws_batch(p(pcm), ctypes.c_size_t(n_clips), ctypes.c_size_t(31 * 1024), p(lg), p(sm), None, ctypes.c_int(0))
    times = []
    for _ in range(reps):
        t0 = time.perf_counter()
        step()
        times.append(time.perf_counter() - t0)
    return times, int(lg.astype(np.int64).sum())


def run(clips_per_proc=128, procs=None, mfcc_only=False, reps=1, seed=20, which=None):
    """returns dict(clips_per_s, frames_per_s, cores, kind, sample, step_s=[...]) ; step time = max over processes"""
    procs = procs or os.cpu_count() or 1
    which = which or kind()      # "reference" = oracle/_ref (firmware-true build), "port" = the C restatement at -O2
    ctx = mp.get_context("spawn")
    with ctx.Pool(procs) as pool:
        res = pool.map(_worker, [(clips_per_proc, seed + i, mfcc_only, reps, which) for i in range(procs)])
    step_s = [max(r[0][k] for r in res) for k in range(reps)]
    total = clips_per_proc * procs
    best = min(step_s)
    return {"clips_per_s": total / best, "frames_per_s": total * 31 / best, "cores": procs, "kind": which,
            "sample": "%d clips x 31 frames (%d per process, %d processes), %s" %
                      (total, clips_per_proc, procs, "q15 MFCC only" if mfcc_only else "q15 MFCC + int8 NNoM net"),
            "step_s": step_s, "clips": total}


if __name__ == "__main__":
    print(run())
