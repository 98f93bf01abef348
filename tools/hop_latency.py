#!/usr/bin/env python3
"""Where a small streaming hop spends its time: K1 (n new frames -> n net-input rows) and K2 (n windows) timed separately and
back to back with CUDA events on one stream, for a few stream counts.  Measurement tool for the <= 0.06 ms / 1,024-stream target.
usage: hop_latency.py [--json out.json]"""
import ctypes, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import numpy as np, torch
    import edison_b200 as ed
    dev = torch.device("cuda", 0)
    ed.init(device=0)
    L, chk = ed.lib(), ed.capi.check
    st = torch.cuda.Stream(device=dev); torch.cuda.set_stream(st)
    sp = ctypes.c_void_p(st.cuda_stream)
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    out = {}
    for n in (32, 1024, 4096, 16384, 100000):
        pcm = (torch.randn(n * 1024, device=dev) * 3000).to(torch.int16)
        rows = torch.empty(n * 13, dtype=torch.int8, device=dev)
        win = torch.randint(-128, 127, (n * 403,), dtype=torch.int8, device=dev)
        sm = torch.empty(n, 10, dtype=torch.int8, device=dev)
        k1 = lambda: chk(L.edison_mfcc_q15_clips_dev(P(pcm), n, 1, 1024, None, P(rows), sp))
        k2 = lambda: chk(L.edison_infer_q7_batch_dev(P(win), n, None, P(sm), sp))
        def both():
            k1(); k2()
        res = {}
        for name, fn in (("k1", k1), ("k2", k2), ("k1+k2", both)):
            for _ in range(20):
                fn()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(201)]
            torch.cuda.synchronize()
            ev[0].record(st)
            for i in range(200):
                fn(); ev[i + 1].record(st)
            torch.cuda.synchronize()
            t = np.array([ev[i].elapsed_time(ev[i + 1]) for i in range(200)])
            res[name] = {"p50_us": round(float(np.percentile(t, 50)) * 1e3, 2), "p99_us": round(float(np.percentile(t, 99)) * 1e3, 2)}
        out[str(n)] = res
        print(n, res, flush=True)
    if "--json" in sys.argv:
        json.dump(out, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)


if __name__ == "__main__":
    main()
