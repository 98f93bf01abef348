#!/usr/bin/env python3
"""A small pass over every kernel of the library for compute-sanitizer (memcheck / racecheck / synccheck / initcheck), SURVEY 5.
Small on purpose (the tools slow kernels down 10-100x): a few thousand frames, both K1 instantiations (quiet and saturating
frames), ragged tails, K2 on full and partial batches (aligned and unaligned input: bulk-copy and byte-load paths), 1-stream
and 37-stream hops in both streaming modes, K3 fp32/fp64, K4, the generic executor.  Results are compared with nothing here --
the parity tests do that -- the point is the sanitizer's report.
usage: compute-sanitizer --tool memcheck python tools/sanitizer_workload.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import common  # noqa: E402  (synthetic inputs only)
import edison_b200 as ed  # noqa: E402


def main():
    ed.init(device=0)
    rng = np.random.default_rng(3)
    # K1: 1 frame, ragged groups, both instantiations (synth_frames mixes quiet, full-scale, DC, Nyquist, impulse frames)
    for nf in (1, 7, 8, 9, 333, 2500):
        ed.mfcc_q15_batch(common.synth_frames(nf, seed=nf))
    ed.mfcc_q15_clips(common.synth_frames(640, seed=2).reshape(-1), 40, 15, 16000)   # 1 s clips: 15 frames at stride 16,000
    # K1 + K2 end to end: partial batches of the net (1, 31, 33, 100 clips)
    for n in (1, 31, 33, 100):
        ed.kws_batch(common.synth_frames(31 * n, seed=10 + n).reshape(-1), n, want_net_in=True)
    # K2 alone from host buffers (unaligned slices -> the byte-load input path on the device side is hit via chunk offsets)
    x = rng.integers(-128, 128, (200, 403), dtype=np.int8)
    for mode in (0, 1, 2):
        ed.set_softmax_mode(mode)
        ed.infer_q7_batch(x)
        ed.infer_q7_batch(x[:1])
        ed.infer_q7_batch(x[3:70])
    ed.set_softmax_mode(ed.SOFTMAX_CMSIS_SIMD)
    # streaming: 1 stream and 37 streams, both modes, past the point where inference starts and every ring wraps
    for n in (1, 37):
        for incremental in (False, True):
            with ed.Streams(n, incremental) as st:
                for h in range(46):
                    st.push_hop(common.synth_frames(n, seed=100 + h))
                st.window(0)
    # K3 float MFCC (numpy-equivalent fp64 / fp32, mfcc.c mode), K4 float CNN, generic executor
    pcm = common.synth_frames(130, seed=4).reshape(-1)
    for dt in (np.float64, np.float32):
        for mode in (0, 1, 2):
            ed.mfcc_float_batch(pcm, mode=mode, dtype=dt)
    with ed.MfccC(13, 1, 512, 8, 0.97) as m:
        m.compute(pcm, hop=256)
    ed.infer_f32_batch(rng.normal(0, 40, (70, 403)).astype(np.float32))
    with ed.Model(os.path.join(ROOT, "tests", "golden", "nnom_kws_example.bin")) as m:
        m.infer(rng.integers(-128, 128, (9, m.in_size), dtype=np.int8))
    ed.shutdown()
    print("sanitizer workload done")


if __name__ == "__main__":
    main()
