#!/usr/bin/env python3
"""Run 60 hops of 100,000 device-resident streams (for an ncu --metrics gpu__time_duration.sum launch list of the two kernels of a hop)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import edison_b200 as ed
ed.init(device=0)
L, chk = ed.lib(), ed.capi.check
ns = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
pcm = (torch.randn(4 * ns * 1024, device="cuda") * 3000).to(torch.int16)
with ed.Streams(ns) as ss:
    for h in range(60):
        chk(L.edison_streams_push_hop_dev(ss.handle, ctypes.c_void_p(pcm[(h % 4) * ns * 1024:].data_ptr()), None, None, None, None))
    torch.cuda.synchronize()
print("done")
