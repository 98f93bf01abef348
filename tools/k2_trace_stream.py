#!/usr/bin/env python3
"""Per-unit clock stamps of K2 inside a streaming hop (block 0, second iteration), for a streaming mode: 0 full window, 1 incremental (row cache).
usage: k2_trace_stream.py [mode] [n_streams]"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
import edison_b200 as ed
from k2_trace import KIND, NU

def main():
    mode = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    ns = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
    ed.init(device=0)
    L, chk = ed.lib(), ed.capi.check
    L.edison_streams_debug_trace.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    pcm = (torch.randn(4 * ns * 1024, device="cuda") * 3000).to(torch.int16)
    tr = torch.zeros(128, dtype=torch.int64, device="cuda")
    with ed.Streams(ns, mode) as ss:
        hop = lambda h: chk(L.edison_streams_push_hop_dev(ss.handle, ctypes.c_void_p(pcm[(h % 4) * ns * 1024:].data_ptr()), None, None, None, None))
        for h in range(40):
            hop(h)
        chk(L.edison_streams_debug_trace(ss.handle, ctypes.c_void_p(tr.data_ptr())))
        hop(40); hop(41)
        torch.cuda.synchronize()
    if mode == 1:   # the incremental kernel's eight-unit schedule; each ' = one batch back
        kind, nu = ["DN''''", "C4'''", "C3(3)''", "C3(4)''", "C2(r4)'", "C1(t2,x0)", "C1(t2,x1)", "C1(t2,x2)"], 8
    else:
        kind, nu = KIND, NU
    t = tr.cpu().numpy()[:4 * nu].reshape(nu, 4)
    t0 = t[:, 0].min()
    print("streaming mode %d, %d streams" % (mode, ns))
    print("%-12s %8s %8s %8s %8s | %7s %7s %7s" % ("unit", "wait", "issue", "ready", "released", "slotwait", "mma", "epilogue"))
    for p in range(nu):
        w, i, r, e = (t[p] - t0).tolist()
        print("%-12s %8d %8d %8d %8d | %7d %7d %7d" % (kind[p], w, i, r, e, i - w, r - i, e - r))
    print("one iteration: last release - first wait = %d cycles" % (t[:, 3].max() - t0))
    if mode == 1:
        f = tr.cpu().numpy()[4 * nu:4 * nu + 2] - t0
        print("dense unit, thread 0: ready %d -> accumulators in registers +%d -> logits / softmax / filter done +%d -> released +%d" % (t[0, 2] - t0, f[0] - (t[0, 2] - t0), f[1] - f[0], (t[0, 3] - t0) - f[1]))

if __name__ == "__main__":
    main()
