#!/usr/bin/env python3
"""Time K2 (int8 NNoM net) of one or more builds of the library on bench.py's 131,072 clips and check logits/softmax against the oracle.
usage: k2_time.py [lib.so ...]   (default: the in-tree library).  Each library runs in its own process."""
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
    net_in = torch.empty(n * 403, dtype=torch.int8, device=dev)
    lg = torch.empty(n, 10, dtype=torch.int8, device=dev); sm = torch.empty(n, 10, dtype=torch.int8, device=dev)
    chk(L.edison_mfcc_q15_clips_dev(P(pcm), n, 31, bench.CLIP, None, P(net_in), sp))
    for _ in range(3):
        chk(L.edison_infer_q7_batch_dev(P(net_in), n, P(lg), P(sm), sp))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record(st)
    for _ in range(20):
        chk(L.edison_infer_q7_batch_dev(P(net_in), n, P(lg), P(sm), sp))
    b.record(st); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    cnt = torch.zeros(10, dtype=torch.int32, device=dev)
    chk(L.edison_infer_q7_counts_dev(P(net_in), n, None, None, P(cnt), sp)); torch.cuda.synchronize()
    hist_ok = bool((cnt.cpu() == torch.bincount(lg.cpu().to(torch.int64).argmax(dim=1), minlength=10).to(torch.int32)).all())
    print("K2 %.4f ms per %d clips | class histogram == argmax counts: %s   (round-1 mma.sync build of the same net: 0.703 ms, profiles/r02b)" % (ms, n, hist_ok), flush=True)
    o = common.Oracle()
    ok = True
    for c0 in range(0, n, n // 6):
        wl, ws = o.forward(net_in[c0 * 403:(c0 + 256) * 403].cpu().numpy())
        ok &= bool((wl == lg[c0:c0 + 256].cpu().numpy()).all() and (ws == sm[c0:c0 + 256].cpu().numpy()).all())
    print("%-30s k2 %.4f ms  %.1f M clips/s  parity %s" % (os.environ.get("EDISON_LIB_TAG", "in-tree"), ms, n / ms / 1e3, "OK" if ok else "MISMATCH"), flush=True)

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
