"""CPU baseline of the KWS hot path on the host cores (TEST / MEASUREMENT INFRASTRUCTURE).

Times the reference's own C built for the host (oracle/_ref, oracle/build_ref.sh), SURVEY 8(d):
  "reference"     B1, firmware-true: firmware/src/audioprocessing.c at -O0 + CMSIS-DSP SIMD branches on host-emulated M4
                  intrinsics, NNoM + CMSIS-NN SIMD branches (the canonical arithmetic: what the GPU path reproduces bit for bit)
  "reference_o3"  B2, the fair CPU figure: the reference's plain-C CMSIS branches at -O3 (AVX2) + NNoM's local backend at -O3
                  (libref_mfcc_c_o3.so + libref_nnom_local_o3.so; +-1 LSB from B1 in the MFCC)
  "port"          the oracle's own C restatement (oracle/libedison_oracle.so, -O2), when oracle/_ref is absent
The reference code is not re-entrant (static buffers), so it is parallelised with one PROCESS per core over disjoint clip
ranges.  Only bench.py's cpu_baseline / --impl reference legs call this.
"""
import ctypes
import multiprocessing as mp
import os
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIBS = {"reference": ("libref_mfcc_simd.so", "libref_nnom_cmsis_simd.so"), "reference_o3": ("libref_mfcc_c_o3.so", "libref_nnom_local_o3.so")}
PORT = os.path.join(HERE, "libedison_oracle.so")
BLOB = os.path.join(os.path.dirname(HERE), "edison_b200", "data", "kws_nnom_q7.bin")


def have(which):
    return which in LIBS and all(os.path.exists(os.path.join(HERE, "_ref", f)) for f in LIBS[which])


def kind():
    return "reference" if have("reference") else "port"


def _synth(n_clips, seed):
    """the distributions the GPU arm sees (bench.py synth_clips_gpu): speech-like sinusoids + noise (D1), every 8th clip
    full-scale uniform (D2: saturating adds), every 16th low level (D3)"""
    rng = np.random.default_rng(seed)
    L = 31 * 1024
    t = np.arange(L, dtype=np.float64)
    out = np.empty((n_clips, L), np.int16)
    for c in range(n_clips):
        if c % 8 == 3:
            x = rng.integers(-32768, 32768, L).astype(np.float64)
        elif c % 16 == 5:
            x = rng.integers(-64, 65, L).astype(np.float64)
        else:
            x = rng.normal(0, 200, L)
            for _ in range(4):
                x += rng.uniform(500, 8000) * np.sin(2 * np.pi * rng.uniform(100, 4000) / 16000 * t + rng.uniform(0, 6.28))
        out[c] = np.clip(np.rint(x), -32768, 32767).astype(np.int16)
    return out.reshape(-1)


def _worker(args):
    n_clips, seed, mfcc_only, reps, which = args
    pcm = _synth(n_clips, seed)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    nf = n_clips * 31
    mf = np.zeros((nf, 32), np.int16)
    lg = np.zeros((n_clips, 10), np.int8)
    sm = np.zeros((n_clips, 10), np.int8)
    if which in LIBS:
        A, N = (ctypes.CDLL(os.path.join(HERE, "_ref", f)) for f in LIBS[which])
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


def run(clips_per_proc=128, procs=None, mfcc_only=False, reps=1, seed=20, which=None, warmup=0):
    """returns dict(clips_per_s, frames_per_s, cores, kind, sample, step_s=[...]); step time = max over processes;
    clips_per_s is the BEST of the `reps` timed repetitions after `warmup` untimed ones"""
    procs = procs or os.cpu_count() or 1
    which = which or kind()
    if which in LIBS and not have(which):
        which = "port"
    ctx = mp.get_context("spawn")
    with ctx.Pool(procs) as pool:
        res = pool.map(_worker, [(clips_per_proc, seed + i, mfcc_only, warmup + reps, which) for i in range(procs)])
    step_s = [max(r[0][k] for r in res) for k in range(warmup, warmup + reps)]
    total = clips_per_proc * procs
    best = min(step_s)
    return {"clips_per_s": total / best, "frames_per_s": total * 31 / best, "cores": procs, "kind": which,
            "sample": "%d clips x 31 frames (%d per process, %d processes; 1/8 full-scale, 1/16 low-level, rest speech-like), %s, best of %d after %d warm-up" %
                      (total, clips_per_proc, procs, "q15 MFCC only" if mfcc_only else "q15 MFCC + int8 NNoM net", reps, warmup),
            "step_s": step_s, "clips": total}


if __name__ == "__main__":
    for w in ("reference", "reference_o3", "port"):
        print(w, run(clips_per_proc=16, reps=2, warmup=1, which=w)["clips_per_s"])
