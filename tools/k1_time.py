#!/usr/bin/env python3
"""Time K1 (q15 MFCC) of one or more builds of the library on bench.py's workload and check each against the oracle.
usage: k1_time.py [lib.so ...]   (default: the in-tree library).  Each library runs in its own process."""
import ctypes, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

def one():
    import numpy as np, torch
    import bench, common
    import edison_b200 as ed
    dev = torch.device("cuda", 0)
    ed.init(device=0)
    L, chk = ed.lib(), ed.capi.check
    st = torch.cuda.Stream(device=dev); torch.cuda.set_stream(st)
    sp = ctypes.c_void_p(st.cuda_stream)
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    n = bench.CLIPS_PER_GPU
    pcm = bench.synth_clips_gpu(torch, n, bench.CLIP, 20, dev)
    if os.environ.get("K1_HALF"):   # every frame below the no-saturation threshold
        pcm = pcm >> 1
    net_in = torch.empty(n * 403, dtype=torch.int8, device=dev)
    for _ in range(3):
        chk(L.edison_mfcc_q15_clips_dev(P(pcm), n, 31, bench.CLIP, None, P(net_in), sp))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record(st)
    for _ in range(10):
        chk(L.edison_mfcc_q15_clips_dev(P(pcm), n, 31, bench.CLIP, None, P(net_in), sp))
    b.record(st); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    # parity: 6 x 64 clips spread over the batch (every distribution of the synthetic mix) against the oracle
    o = common.Oracle()
    ok = True
    for c0 in range(0, n, n // 6):
        sub = pcm[c0 * bench.CLIP:(c0 + 64) * bench.CLIP].cpu().numpy()
        want = o.net_input(o.mfcc(sub.reshape(-1, 1024))).reshape(-1)
        got = net_in[c0 * 403:(c0 + 64) * 403].cpu().numpy()
        ok &= bool((want == got).all())
    print("%-40s k1 %.3f ms  %.1f M frames/s  parity %s" % (os.environ.get("EDISON_LIB_TAG", "in-tree"), ms, n * 31 / ms / 1e3, "OK" if ok else "MISMATCH"), flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--one":
        one()
    else:
        libs = sys.argv[1:] or [None]
        dst = os.path.join(ROOT, "edison_b200", "libedison_b200.so")
        keep = open(dst, "rb").read()
        try:
            for lib in libs:
                env = dict(os.environ)
                if lib:
                    open(dst, "wb").write(open(lib, "rb").read()); env["EDISON_LIB_TAG"] = os.path.basename(lib)
                subprocess.call([sys.executable, os.path.abspath(__file__), "--one"], env=env)
        finally:
            open(dst, "wb").write(keep)
