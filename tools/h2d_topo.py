"""Pinned host -> device copy rates per SET of GPUs, from one process (is the e2e ceiling at 8 GPUs a shared
uplink or a host-DRAM limit?).  For every set, each GPU of the set copies its own pinned 1 GiB buffer `reps` times on its own
stream, all started together; the rate is the bytes of all GPUs / the wall time until the last one finished.

    python tools/h2d_topo.py --out gpurun_out/h2d_topo.json        (on a box with N visible GPUs)

Also records `nvidia-smi topo -m`.  Measurement tool only; nothing in the library uses it."""
import argparse
import json
import subprocess
import time

import torch


def rate(devs, bufs_h, bufs_d, reps):
    streams = [torch.cuda.Stream(device=d) for d in devs]
    for d in devs:
        torch.cuda.synchronize(d)
    t0 = time.perf_counter()
    for _ in range(reps):
        for d, s in zip(devs, streams):
            with torch.cuda.stream(s):
                bufs_d[d].copy_(bufs_h[d], non_blocking=True)
    for d in devs:
        torch.cuda.synchronize(d)
    dt = time.perf_counter() - t0
    return len(devs) * reps * bufs_h[devs[0]].numel() / dt / 1e9


def wc_rate(dst, size, reps):
    """device 0, one stream: cudaHostAlloc(default) vs cudaHostAlloc(cudaHostAllocWriteCombined) as the source of the same copy"""
    import ctypes
    rt = ctypes.CDLL("libcudart.so.12")
    rt.cudaHostAlloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t, ctypes.c_uint]
    rt.cudaMemcpy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
    rt.cudaFreeHost.argtypes = [ctypes.c_void_p]
    res = {}
    torch.cuda.set_device(0)
    for name, flag in (("pinned", 0), ("pinned_write_combined", 4)):
        p = ctypes.c_void_p()
        if rt.cudaHostAlloc(ctypes.byref(p), size, flag) != 0:
            res[name] = None
            continue
        ctypes.memset(p, 1, size)
        best = 0.0
        for _ in range(3):
            torch.cuda.synchronize(0)
            t0 = time.perf_counter()
            for _ in range(reps):
                rt.cudaMemcpy(ctypes.c_void_p(dst.data_ptr()), p, size, 1)
            torch.cuda.synchronize(0)
            best = max(best, reps * size / (time.perf_counter() - t0) / 1e9)
        res[name] = round(best, 1)
        rt.cudaFreeHost(p)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/h2d_topo.json")
    ap.add_argument("--mib", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=8)
    a = ap.parse_args()
    n = torch.cuda.device_count()
    size = a.mib << 20
    bufs_h = {d: torch.empty(size, dtype=torch.uint8).pin_memory() for d in range(n)}
    bufs_d = {d: torch.empty(size, dtype=torch.uint8, device="cuda:%d" % d) for d in range(n)}
    sets = [[0]]
    if n >= 2:
        sets += [[0, d] for d in range(1, n)]
    if n >= 4:
        sets += [[0, 1, 2, 3]]
    if n >= 8:
        sets += [[4, 5, 6, 7], [0, 2, 4, 6], [0, 1, 4, 5], list(range(8))]
    out = {"gib_per_copy": a.mib / 1024, "reps": a.reps, "unit": "GB/s", "sets": {}}
    for s in sets:
        rate(s, bufs_h, bufs_d, 1)   # warm-up
        out["sets"][",".join(map(str, s))] = round(max(rate(s, bufs_h, bufs_d, a.reps) for _ in range(3)), 1)
    out["write_combined"] = wc_rate(bufs_d[0], size, a.reps)
    try:
        out["topo"] = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=60).stdout
    except Exception as e:   # noqa: BLE001
        out["topo"] = "unavailable: %s" % e
    with open(a.out, "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps(out["sets"]))


if __name__ == "__main__":
    main()
