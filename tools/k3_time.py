#!/usr/bin/env python3
"""Time K3 (float MFCC, fp32 and fp64) of one or more builds of the library on bench.py's config #2 frames.
usage: k3_time.py [lib.so ...]   (default: the in-tree library).  Each library runs in its own process."""
import ctypes, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

def one():
    import torch
    import bench
    import edison_b200 as ed
    dev = torch.device("cuda", 0)
    ed.init(device=0)
    L, chk = ed.lib(), ed.capi.check
    st = torch.cuda.Stream(device=dev); torch.cuda.set_stream(st)
    sp = ctypes.c_void_p(st.cuda_stream)
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    pcm = bench.synth_clips_gpu(torch, bench.CFG2_CLIPS, bench.CFG2_STRIDE, 20, dev)
    nfr = bench.CFG2_CLIPS * bench.CFG2_FPC
    out = []
    for name, fn, tdt in (("f32", L.edison_mfcc_f32_clips_dev, torch.float32), ("f64", L.edison_mfcc_f64_clips_dev, torch.float64)):
        o = torch.empty(nfr, 32, dtype=tdt, device=dev)
        for _ in range(3):
            chk(fn(P(pcm), bench.CFG2_CLIPS, bench.CFG2_FPC, bench.CFG2_STRIDE, P(o), 0, sp))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record(st)
        for _ in range(10):
            chk(fn(P(pcm), bench.CFG2_CLIPS, bench.CFG2_FPC, bench.CFG2_STRIDE, P(o), 0, sp))
        b.record(st); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 10
        out.append("%s %.3f ms %.0f M frames/s" % (name, ms, nfr / ms / 1e3))
    print("%-30s %s" % (os.environ.get("EDISON_LIB_TAG", "in-tree"), "   ".join(out)), flush=True)

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
