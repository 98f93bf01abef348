#!/usr/bin/env python3
"""Device-resident hop time of the streaming modes (0 = full window, 1 = incremental: the bottom row of every layer + row cache) for a few stream
counts: p50 / p99 over 300 hops, CUDA events on the launching stream.  usage: stream_modes_time.py [n ...]"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import edison_b200 as ed
ed.init(device=0)
L, chk = ed.lib(), ed.capi.check
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
sp = ctypes.c_void_p(st.cuda_stream)
for ns in [int(a) for a in sys.argv[1:]] or [1024, 16384, 100000]:
    pcm = (torch.randn(4 * ns * 1024, device="cuda") * 3000).to(torch.int16)
    for mode in (0, 1):
        with ed.Streams(ns, mode) as ss:
            hop = lambda h: chk(L.edison_streams_push_hop_dev(ss.handle, ctypes.c_void_p(pcm[(h % 4) * ns * 1024:].data_ptr()), None, None, None, sp))
            for h in range(40):
                hop(h)
            torch.cuda.synchronize()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(301)]
            ev[0].record(st)
            for h in range(300):
                hop(40 + h); ev[h + 1].record(st)
            torch.cuda.synchronize()
            t = np.array([ev[i].elapsed_time(ev[i + 1]) for i in range(300)])
            print("%7d streams, mode %d: hop p50 %.1f us, p99 %.1f us" % (ns, mode, np.percentile(t, 50) * 1e3, np.percentile(t, 99) * 1e3), flush=True)
