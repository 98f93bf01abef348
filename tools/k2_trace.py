#!/usr/bin/env python3
"""Timeline of one steady-state iteration of K2 (block 0, second batch): per schedule unit, when the issuing thread started waiting
for the TMEM slot, when it got it, when the accumulators were ready for the epilogue, and when the epilogue released the slot.
Cycles are SM clocks relative to the first stamp.  usage: k2_trace.py [n_clips]"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import edison_b200 as ed

KIND = ["C1(t0,x0)", "C3(0)'", "C1(t0,x1)", "C3(1)'", "C1(t0,x2)", "C3(2)'", "C1(t1,x0)", "C3(3)'", "C1(t1,x1)", "C3(4)'", "C1(t1,x2)", "C1(t2,x0)",
        "C1(t2,x1)", "C4'", "C1(t2,x2)", "C2(r0)", "C2(r1)", "DN'+T", "C2(r2)", "C2(r3)", "C2(r4)"]   # ' = previous batch
FINE = 17
NU = len(KIND)

def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
    dev = torch.device("cuda", 0)
    ed.init(device=0)
    L, chk = ed.lib(), ed.capi.check
    L.edison_debug_k2_trace_dev.argtypes = [ctypes.c_void_p, ctypes.c_size_t] + [ctypes.c_void_p] * 4
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    g = torch.Generator(device=dev); g.manual_seed(1)
    net_in = torch.randint(-128, 128, (n, 403), dtype=torch.int8, device=dev, generator=g)
    lg = torch.empty(n, 10, dtype=torch.int8, device=dev); sm = torch.empty_like(lg)
    tr = torch.zeros(4 * NU + 4, dtype=torch.int64, device=dev)   # [4 NU .. 4 NU + 2]: fine stamps of unit FINE
    for _ in range(2):
        chk(L.edison_debug_k2_trace_dev(P(net_in), n, P(lg), P(sm), P(tr), None)); chk(L.edison_sync(None))
    t = tr.cpu().numpy()[:4 * NU].reshape(NU, 4)
    t0 = t[:, 0].min()
    print("%-12s %8s %8s %8s %8s | %7s %7s %7s" % ("unit", "wait", "issue", "ready", "released", "slotwait", "mma", "epilogue"))
    for p in range(NU):
        w, i, r, e = (t[p] - t0).tolist()
        print("%-12s %8d %8d %8d %8d | %7d %7d %7d" % (KIND[p], w, i, r, e, i - w, r - i, e - r))
    print("one iteration: last release - first wait = %d cycles" % (t[:, 3].max() - t0))
    f = tr.cpu().numpy()[4 * NU:4 * NU + 3] - t0
    r8, e8 = (t[FINE, 2] - t0), (t[FINE, 3] - t0)
    print("team thread 0, unit C2(r0): ready %d -> accumulators in registers +%d -> stored +%d -> proxy fence +%d -> arrived +%d" %
          (r8, f[0] - r8, f[1] - f[0], f[2] - f[1], e8 - f[2]))

if __name__ == "__main__":
    main()
