#!/usr/bin/env python3
"""bench.py -- KWS hot path throughput on B200 (contract: see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step = one pass of the hot path (q15 MFCC front-end K1 + int8 NNoM net K2) over this rank's shard
of synthetic clips: BASELINE.json config #3 (2^20 clips over 8 GPUs) = 131,072 clips x 31 frames of
1024 int16 samples per GPU (8.3 GB, far larger than L2), weak scaling.  `value` times the kernels
with the PCM resident in HBM; `e2e` times edison_kws_batch() -- the C-ABI call a user makes -- on
pinned HOST buffers, copies included.  The same line carries MFCC-only frames/s on config #2
(65,536 one-second clips, 15 frames each), the roofline of the dominant kernel and the CPU
reference timed on this box's cores.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CLIPS_PER_GPU = 131072          # config #3: 2^20 clips / 8 GPUs
CFG2_CLIPS, CFG2_FPC, CFG2_STRIDE = 65536, 15, 16000
E2E_CLIPS = 32768               # host-buffer leg: 2.08 GB of pinned PCM per step
CLIP = 31 * 1024
K1_BYTES_PER_FRAME = 2048 + 13  # algorithmic: PCM read + int8 net-input row written
N_STREAMS = 100000              # config #5


def k1_ncu():
    """dram bytes per launch and warp instructions per frame of K1 from the committed ncu summary (profiles/k1_ncu.json, written by
    tools/ncu_summary.py --json from an `ncu --set full` capture of this bench's K1 launch): (traffic, inst/frame, source)"""
    try:
        with open(os.path.join(ROOT, "profiles", "k1_ncu.json")) as fh:
            d = json.load(fh)
        return float(d["dram_bytes_per_launch"]) * (CLIPS_PER_GPU * 31) / float(d["frames"]), float(d["warp_inst_per_frame"]), d.get("source")
    except Exception:
        return None, None, None


def int8_peak():
    """dense int8 tensor rate measured on this pool's B200 with tcgen05.mma.kind::i8 from shared memory on all SMs
    (tools/ubench/umma_probe.cu -> profiles/r02a_umma_probe.json); 2 x the measured bf16 rate only if that file is missing"""
    try:
        with open(os.path.join(ROOT, "profiles", "r02a_umma_probe.json")) as fh:
            return float(json.load(fh)["int8_tops_measured"]), "measured (tools/ubench/umma_probe.cu, profiles/r02a_umma_probe.json)"
    except Exception:
        bf = peaks_bf16()
        return 2 * bf[0], "2 x bf16 " + bf[1]


def bind_to_gpu_numa(index):
    """Pin this process to the CPUs NVML reports as local to GPU `index`, so that pinned host buffers (first touch) and the
    copy-issuing thread sit on the GPU's own NUMA node.  Returns the number of CPUs bound to, or None when unavailable."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:   # CUDA's device order need not be NVML's: go through the PCI address
            import torch
            pr = torch.cuda.get_device_properties(index)
            h = pynvml.nvmlDeviceGetHandleByPciBusId(("%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)).encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(index)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * i + b for i, wd in enumerate(words) for b in range(64) if (wd >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def peaks_bf16():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["bf16_tflops"]), "measured"
    except Exception:
        return 2250.0, "nominal"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return json.load(fh)["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons while the timed region runs (NVML every 5 ms; nvidia-smi as a fallback)"""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.mhz, self.reasons, self.max_mhz, self.stop_flag = index, [], set(), None, False
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def run(self):
        n = self.nvml
        while not self.stop_flag:
            try:
                if n:
                    self.mhz.append(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM))
                    r = n.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
                        else n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for name, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20),
                                      ("hw_thermal_slowdown", 0x40)):
                        if r & bit:
                            self.reasons.add(name)
                    time.sleep(0.005)
                else:
                    out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout.strip()
                    c = [v.strip() for v in out.split(",")]
                    if c and c[0].isdigit():
                        self.mhz.append(int(c[0]))
                        self.max_mhz = int(c[1])
                        for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], c[2:6]):
                            if v.lower().startswith("active"):
                                self.reasons.add(name)
                    time.sleep(0.05)
            except Exception:
                time.sleep(0.05)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        m = sorted(self.mhz)
        return {"sm_mhz": m[len(m) // 2] if m else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(m), "source": "nvml" if self.nvml else "nvidia-smi"}


def synth_clips_gpu(torch, n_clips, stride, seed, dev):
    """int16 [n_clips * stride] on the device: speech-like sinusoids + noise (D1), every 8th clip full-scale
    uniform (D2, takes the saturating path), every 16th low level (D3).  Deterministic per seed."""
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    out = torch.empty(n_clips * stride, dtype=torch.int16, device=dev)
    view = out.view(n_clips, stride)
    t = torch.arange(stride, device=dev, dtype=torch.float32)
    step = 4096
    for c0 in range(0, n_clips, step):
        nc = min(step, n_clips - c0)
        x = torch.randn(nc, stride, generator=g, device=dev) * 200.0
        for _ in range(4):
            amp = torch.rand(nc, 1, generator=g, device=dev) * 7500 + 500
            f = torch.rand(nc, 1, generator=g, device=dev) * 3900 + 100
            ph = torch.rand(nc, 1, generator=g, device=dev) * 6.28
            x += amp * torch.sin(2 * 3.14159265 * f / 16000.0 * t + ph)
        idx = torch.arange(c0, c0 + nc, device=dev)
        full = (idx % 8 == 3).view(-1, 1)
        low = (idx % 16 == 5).view(-1, 1)
        u = torch.rand(nc, stride, generator=g, device=dev)
        x = torch.where(full, u * 65535.0 - 32768.0, x)
        x = torch.where(low, u * 128.0 - 64.0, x)
        view[c0:c0 + nc] = x.round().clamp(-32768, 32767).to(torch.int16)
        del x, u
    return out


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation on the host cores (rank 0 only).  The line's value is the FASTEST build of
    the reference's sources, so that the ratio the driver computes is not flattered by a slow denominator: B2, the reference's plain-C CMSIS
    path at -O3 (AVX2) + NNoM local backend, when it was built; else B1, the firmware-true build (audioprocessing.c -O0 + CMSIS SIMD branches on
    emulated M4 intrinsics: the arithmetic the GPU arm reproduces bit for bit).  The other builds and the oracle's plain-C port are timed
    beside it and reported in the same line."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import cpu_baseline
    cores = os.cpu_count() or 1
    per = 32
    builds = {"reference": "B1 firmware-true: audioprocessing.c -O0 + CMSIS SIMD branches on emulated M4 intrinsics, NNoM + CMSIS-NN",
              "reference_o3": "B2: the reference's plain-C CMSIS branches at -O3 -march=x86-64-v3 + NNoM local backend at -O3 (SURVEY 8(d) fair CPU figure)"}
    main = "reference_o3" if cpu_baseline.have("reference_o3") else "reference"
    res = cpu_baseline.run(clips_per_proc=per, procs=cores, mfcc_only=False, reps=args.steps, warmup=args.warmup, which=main)
    steps = res["step_s"]
    ms = 1e3 * sum(steps) / len(steps)
    val = res["clips"] / (ms / 1e3)
    others = {}
    for which in ("reference", "reference_o3", "port"):
        if which == main or (which != "port" and not cpu_baseline.have(which)):
            continue
        r2 = cpu_baseline.run(clips_per_proc=per, procs=cores, reps=3, warmup=1, which=which)
        others[r2["kind"]] = {"value": r2["clips_per_s"], "unit": "clips/s", "cores": r2["cores"], "sample": r2["sample"],
                              "build": builds.get(which, "oracle/libedison_oracle.so (plain-C restatement, -O2)")}
    line = {"impl": "reference", "metric": "kws_clips_per_s", "value": val, "unit": "clips/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "q15/int8", "data": "synthetic",
            "config": {"workload": "config#3 shard: clips x 31 frames x 1024 int16, q15 MFCC + int8 NNoM net",
                       "sample": res["sample"]},
            "cpu_baseline": {"value": val, "unit": "clips/s", "cores": cores, "kind": "reference" if res["kind"].startswith("reference") else "port",
                             "sample": res["sample"], "build": builds.get(main, main)},
            "cpu_baseline_other_builds": others,
            "e2e": {"value": val, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "mfcc_frames_per_s": val * 31, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_ours(args, rank, world, local_rank):
    import torch
    import edison_b200 as ed
    from edison_b200 import shard
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = bind_to_gpu_numa(local_rank)   # before any pinned allocation: the e2e leg streams host memory over PCIe
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    ed.init(device=local_rank)
    L = ed.lib()
    chk = ed.capi.check
    st = torch.cuda.Stream(device=dev)   # a real (non-NULL) stream: the library launches on the handle it is given
    torch.cuda.set_stream(st)
    sp = ctypes.c_void_p(st.cuda_stream)
    P = lambda t: ctypes.c_void_p(t.data_ptr())

    n = CLIPS_PER_GPU
    pcm = synth_clips_gpu(torch, n, CLIP, 20 + rank, dev)
    net_in = torch.empty(n * 403, dtype=torch.int8, device=dev)
    logits = torch.empty(n, 10, dtype=torch.int8, device=dev)
    softmax = torch.empty(n, 10, dtype=torch.int8, device=dev)
    gathered = torch.empty(world * n, 10, dtype=torch.int8, device=dev) if world > 1 else None
    counts = torch.zeros(10, dtype=torch.int32, device=dev)     # class histogram, accumulated by K2's dense epilogue on the device
    total_counts = None

    def step(ev=None):
        if ev: ev[0].record(st)
        chk(L.edison_mfcc_q15_clips_dev(P(pcm), n, 31, CLIP, None, P(net_in), sp))
        if ev: ev[1].record(st)
        chk(L.edison_infer_q7_counts_dev(P(net_in), n, P(logits), P(softmax), P(counts), sp))
        if ev: ev[2].record(st)

    def final_exchange():
        """the path's only exchange, once after the last batch: gather of the logits, sum of the class counts"""
        if world > 1:
            dist.all_gather_into_tensor(gathered, logits)
            return shard.reduce_counts(dist, counts)
        return counts.to(torch.int64)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    final_exchange()
    counts.zero_()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    l0 = ed.launch_count()
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_begin.record(st)
    for k in range(args.steps):
        step(evs[k])
    t_comm = torch.cuda.Event(enable_timing=True)
    t_comm.record(st)
    total_counts = final_exchange()          # inside the timed region: the job is K batches, then the exchange
    t_end.record(st)
    barrier()
    launches = ed.launch_count() - l0
    total_ms = t_begin.elapsed_time(t_end)
    comm_ms = t_comm.elapsed_time(t_end)
    k1_ms = sum(e[0].elapsed_time(e[1]) for e in evs) / args.steps
    k2_ms = sum(e[1].elapsed_time(e[2]) for e in evs) / args.steps
    clocks = sampler.summary()
    if world > 1:
        t = torch.tensor([total_ms, k1_ms, k2_ms, comm_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, k1_ms, k2_ms, comm_ms = t.tolist()
        # on the ranks that finish their batches early the exchange also contains the wait for the slowest rank; the smallest value over the
        # ranks is the exchange itself
        tmin = torch.tensor([t_comm.elapsed_time(t_end)], device=dev, dtype=torch.float64)
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        comm_ms_min = tmin.item()
    else:
        comm_ms_min = comm_ms
    ms_per_step = total_ms / args.steps
    value = world * n / (ms_per_step / 1e3)

    # ---- what the timed region computed, checked AFTER it: sampled clips of the last step against the CPU oracle (bit-exact),
    # and the device-side class histogram against an argmax of the logits it was built from
    parity = None
    if rank == 0:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import numpy as np
        import common
        pick = np.linspace(0, n - 1, 256).astype(np.int64)
        host_pcm = torch.stack([pcm[int(c) * CLIP:(int(c) + 1) * CLIP] for c in pick]).cpu().numpy()
        want_lg, want_sm, want_ni = common.Oracle().kws(host_pcm.reshape(-1), len(pick))
        idx = torch.from_numpy(pick).to(dev)
        got_lg, got_sm = logits[idx].cpu().numpy(), softmax[idx].cpu().numpy()
        got_ni = net_in.view(n, 403)[idx].cpu().numpy()
        hist_ok = bool((counts.to(torch.int64) == args.steps * torch.bincount(logits.to(torch.int64).argmax(dim=1), minlength=10)).all().item())
        parity = {"clips_checked": int(len(pick)), "net_input_exact": bool((got_ni == want_ni).all()), "logits_exact": bool((got_lg == want_lg).all()),
                  "softmax_exact": bool((got_sm == want_sm).all()), "class_histogram_matches_logits": hist_ok,
                  "class_counts_total": int(total_counts.sum().item()), "against": "oracle/libedison_oracle.so (pinned to oracle/_ref)"}
        parity["ok"] = parity["net_input_exact"] and parity["logits_exact"] and parity["softmax_exact"] and hist_ok and \
            parity["class_counts_total"] == args.steps * world * n

    # ---- MFCC-only on config #2 (65,536 one-second clips -> 15 frames each), device resident
    pcm2 = pcm[: CFG2_CLIPS * CFG2_STRIDE]
    mf2 = torch.empty(CFG2_CLIPS * CFG2_FPC, 32, dtype=torch.int16, device=dev)
    for _ in range(max(args.warmup, 3)):
        chk(L.edison_mfcc_q15_clips_dev(P(pcm2), CFG2_CLIPS, CFG2_FPC, CFG2_STRIDE, P(mf2), None, sp))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record(st)
    for _ in range(args.steps):
        chk(L.edison_mfcc_q15_clips_dev(P(pcm2), CFG2_CLIPS, CFG2_FPC, CFG2_STRIDE, P(mf2), None, sp))
    b.record(st)
    torch.cuda.synchronize()
    cfg2_ms = a.elapsed_time(b) / args.steps
    cfg2_fps = CFG2_CLIPS * CFG2_FPC / (cfg2_ms / 1e3)

    # ---- float MFCC (K3; config #1's arithmetic at config #2's size), device resident: fp64 parity kernel and fp32 fast kernel
    float_stats = {}
    for name, fn, tdt in (("f64", L.edison_mfcc_f64_clips_dev, torch.float64), ("f32", L.edison_mfcc_f32_clips_dev, torch.float32)):
        outf = torch.empty(CFG2_CLIPS * CFG2_FPC, 32, dtype=tdt, device=dev)
        for _ in range(3):
            chk(fn(P(pcm2), CFG2_CLIPS, CFG2_FPC, CFG2_STRIDE, P(outf), 0, sp))
        torch.cuda.synchronize()
        a.record(st)
        for _ in range(args.steps):
            chk(fn(P(pcm2), CFG2_CLIPS, CFG2_FPC, CFG2_STRIDE, P(outf), 0, sp))
        b.record(st)
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / args.steps
        nfr = CFG2_CLIPS * CFG2_FPC
        float_stats[name] = {"frames_per_s": nfr / (ms / 1e3), "ms": ms,
                             "hbm_gbs": nfr * (2048 + 32 * outf.element_size()) / (ms / 1e3) / 1e9}
        del outf

    # ---- float CNN (K4; config #4: 262,144 clips x 403 float32 = 422.6 MB in, device resident)
    n4 = 262144
    x4 = net_in[: n4 // 2 * 403].to(torch.float32).repeat(2)     # int8 MFCC rows of the KWS step, as floats
    lg4 = torch.empty(n4, 10, dtype=torch.float32, device=dev)
    sm4 = torch.empty(n4, 10, dtype=torch.float32, device=dev)
    for _ in range(3):
        chk(L.edison_infer_f32_batch_dev(P(x4), n4, P(lg4), P(sm4), sp))
    torch.cuda.synchronize()
    a.record(st)
    for _ in range(args.steps):
        chk(L.edison_infer_f32_batch_dev(P(x4), n4, P(lg4), P(sm4), sp))
    b.record(st)
    torch.cuda.synchronize()
    cfg4_ms = a.elapsed_time(b) / args.steps
    cfg4 = {"workload": "config#4: float CNN (Cube.AI kws_conv) on 262,144 clips x 403 float32, 1 GPU", "ms": cfg4_ms,
            "clips_per_s": n4 / (cfg4_ms / 1e3), "fp32_tflops": n4 * 1569504 / (cfg4_ms / 1e3) / 1e12}
    del x4, lg4, sm4

    # ---- generic NNoM executor: the upstream keyword_spotting example of the reference tree (63 x 12 x 1 input,
    # 4 conv + dense, 35 classes, 2.1 M MAC per input) on 65,536 device-resident inputs; tensor-core nodes vs the one-thread-per-output path
    generic = None
    blob = os.path.join(ROOT, "tests", "golden", "nnom_kws_example.bin")
    if rank == 0 and os.path.exists(blob):
        ng = 65536
        with ed.Model(blob) as gm:
            xg = torch.randint(-128, 128, (ng, gm.in_size), dtype=torch.int8, device=dev)
            lgg = torch.empty(ng, gm.n_out, dtype=torch.int8, device=dev)
            smg = torch.empty(ng, gm.n_out, dtype=torch.int8, device=dev)
            outs = {}
            for name, naive, reps in (("tensor_core_nodes", False, 10), ("one_thread_per_output", True, 3)):
                gm.set_naive(naive)
                for _ in range(2):
                    chk(L.edison_model_infer(gm._h, P(xg), ng, P(lgg), P(smg)))       # synchronous (device pointers: no copies)
                t0 = time.perf_counter()
                for _ in range(reps):
                    chk(L.edison_model_infer(gm._h, P(xg), ng, P(lgg), P(smg)))
                outs[name] = ((time.perf_counter() - t0) / reps, lgg.clone())
        generic = {"workload": "generic executor: upstream NNoM keyword_spotting example (tests/golden/nnom_kws_example.bin), 65,536 inputs x 756 int8, device resident",
                   "inputs_per_s": ng / outs["tensor_core_nodes"][0], "ms": outs["tensor_core_nodes"][0] * 1e3,
                   "int8_tops": ng * 2 * 2098240 / outs["tensor_core_nodes"][0] / 1e12,
                   "one_thread_per_output_ms": outs["one_thread_per_output"][0] * 1e3,
                   "speedup_over_one_thread_per_output": outs["one_thread_per_output"][0] / outs["tensor_core_nodes"][0],
                   "same_logits_both_paths": bool(torch.equal(outs["tensor_core_nodes"][1], outs["one_thread_per_output"][1]))}
        del xg, lgg, smg, outs

    # ---- e2e: edison_kws_batch() on pinned HOST buffers, H2D + D2H inside the timed region
    ne = E2E_CLIPS
    h_pcm = torch.empty(ne * CLIP, dtype=torch.int16).pin_memory()
    h_pcm.copy_(pcm[: ne * CLIP])
    h_logits = torch.empty(ne, 10, dtype=torch.int8).pin_memory()
    h_softmax = torch.empty(ne, 10, dtype=torch.int8).pin_memory()
    for _ in range(2):
        chk(L.edison_kws_batch(P(h_pcm), ne, CLIP, P(h_logits), P(h_softmax), None))
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        chk(L.edison_kws_batch(P(h_pcm), ne, CLIP, P(h_logits), P(h_softmax), None))
    e2e_s = (time.perf_counter() - t0) / args.steps
    if world > 1:
        t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = t.item()
    e2e_val = world * ne / e2e_s
    # what bounds this leg: the pinned host -> device copy rate of this box, measured on the same buffer right here
    hb0, hb1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    d_tmp = torch.empty(ne * CLIP, dtype=torch.int16, device=dev)
    d_tmp.copy_(h_pcm, non_blocking=True)
    barrier()
    hb0.record(st)
    for _ in range(3):
        d_tmp.copy_(h_pcm, non_blocking=True)
    hb1.record(st)
    torch.cuda.synchronize()
    h2d_gbs = 3 * ne * CLIP * 2 / (hb0.elapsed_time(hb1) / 1e3) / 1e9
    del d_tmp
    same = bool((h_logits.to(dev) == logits[:ne]).all().item())

    # ---- streaming (config #5): N_STREAMS concurrent streams, one hop = 1024 new samples per stream.
    # A hop is two launches: K1 (one frame per stream, rows stored straight into the ring windows) and K2 (full window per stream,
    # filter / argmax / threshold fused).  Device-resident hops are timed with CUDA events; the ingest leg takes the hop PCM from
    # pinned host memory through the pipelined submit / collect calls (copy of hop h + 1 under the kernels of hop h).
    stream_stats = None
    if world == 1:
        ns = N_STREAMS
        d_sm = torch.empty(ns, 10, dtype=torch.int8, device=dev)
        d_fl = torch.empty(ns, 10, dtype=torch.float32, device=dev)
        d_dt = torch.empty(ns, dtype=torch.int32, device=dev)
        pct = lambda v, q: v[min(len(v) - 1, int(len(v) * q))]
        with ed.Streams(ns) as ss:
            hop = lambda h: chk(L.edison_streams_push_hop_dev(ss.handle, P(pcm[(h % 36) * ns * 1024:]), None, None, None, sp))
            for h in range(34):   # fill the 31-row windows, then two full hops as warm-up
                hop(h)
            torch.cuda.synchronize()
            n_hops = 2000
            hev = [torch.cuda.Event(enable_timing=True) for _ in range(n_hops + 1)]
            hev[0].record(st)
            for h in range(n_hops):
                hop(34 + h)
                hev[h + 1].record(st)
            torch.cuda.synchronize()
            lat = sorted(hev[i].elapsed_time(hev[i + 1]) for i in range(n_hops))
            p50, p99 = pct(lat, 0.5), pct(lat, 0.99)
        with ed.Streams(ns, False) as ss:   # the firmware's own scheme: the whole window through the net every hop (no row cache)
            hopf = lambda h: chk(L.edison_streams_push_hop_dev(ss.handle, P(pcm[(h % 36) * ns * 1024:]), None, None, None, sp))
            for h in range(34):
                hopf(h)
            torch.cuda.synchronize()
            fev = [torch.cuda.Event(enable_timing=True) for _ in range(301)]
            fev[0].record(st)
            for h in range(300):
                hopf(34 + h)
                fev[h + 1].record(st)
            torch.cuda.synchronize()
            latf = sorted(fev[i].elapsed_time(fev[i + 1]) for i in range(300))
            full_window = {"hop_ms_p50": pct(latf, 0.5), "hop_ms_p99": pct(latf, 0.99), "hops_timed": 300,
                           "note": "edison_streams_set_incremental(s, 0): every hop recomputes the whole 31-row window, as the firmware does"}
        with ed.Streams(ns) as ss:
            h_hop = [torch.empty(ns * 1024, dtype=torch.int16).pin_memory() for _ in range(2)]
            for b_ in h_hop:
                b_.copy_(pcm[: ns * 1024])
            h_sm = torch.empty(ns, 10, dtype=torch.int8).pin_memory()
            h_fl = torch.empty(ns, 10, dtype=torch.float32).pin_memory()
            h_dt = torch.empty(ns, dtype=torch.int32).pin_memory()
            n_ing = 500
            chk(L.edison_streams_submit_hop(ss.handle, P(h_hop[0])))
            stamps = []
            for h in range(1, 34 + n_ing):
                chk(L.edison_streams_submit_hop(ss.handle, P(h_hop[h & 1])))
                chk(L.edison_streams_collect_hop(ss.handle, P(h_sm), P(h_fl), P(h_dt)))
                stamps.append(time.perf_counter())
            chk(L.edison_streams_collect_hop(ss.handle, P(h_sm), P(h_fl), P(h_dt)))
            lat_e2e = sorted((stamps[i + 1] - stamps[i]) * 1e3 for i in range(33, len(stamps) - 1))
        small = {}
        with ed.Streams(1024) as s2:
            for h in range(34):
                chk(L.edison_streams_push_hop_dev(s2.handle, P(pcm[(h % 36) * 1024 * 1024:]), None, None, None, sp))
            torch.cuda.synchronize()
            lat2 = []
            for h in range(500):
                t0 = time.perf_counter()
                chk(L.edison_streams_push_hop_dev(s2.handle, P(pcm[(h % 36) * 1024 * 1024:]), None, None, None, sp))
                torch.cuda.synchronize()
                lat2.append((time.perf_counter() - t0) * 1e3)
            lat2.sort()
            ev2 = [torch.cuda.Event(enable_timing=True) for _ in range(501)]
            ev2[0].record(st)
            for h in range(500):
                chk(L.edison_streams_push_hop_dev(s2.handle, P(pcm[(h % 36) * 1024 * 1024:]), None, None, None, sp))
                ev2[h + 1].record(st)
            torch.cuda.synchronize()
            dev2 = sorted(ev2[i].elapsed_time(ev2[i + 1]) for i in range(500))
            small = {"streams": 1024, "hop_ms_p50": pct(lat2, 0.5), "hop_ms_p99": pct(lat2, 0.99), "hops_timed": 500,
                     "note": "host wall clock, launch to completion, one hop at a time",
                     "device_hop_ms_p50": pct(dev2, 0.5), "device_hop_ms_p99": pct(dev2, 0.99), "launches_per_hop": 2}
        stream_stats = {"workload": "config#5: %d concurrent streams, hop 1024 samples (64 ms), window 31 rows, device resident" % ns,
                        "mode": "incremental (default): K2 computes the bottom row of every layer per hop, older rows from its row cache in HBM; bit-identical to the full window",
                        "hop_ms_p50": p50, "hop_ms_p99": p99, "hops_timed": n_hops, "launches_per_hop": 2, "full_window": full_window,
                        "stream_hops_per_s": ns / (p50 / 1e3), "realtime_streams_per_gpu": int(ns * 64.0 / p99),
                        "ingest_hop_ms_p50": pct(lat_e2e, 0.5), "ingest_hop_ms_p99": pct(lat_e2e, 0.99), "ingest_hops_timed": len(lat_e2e),
                        "ingest_bytes_per_hop": ns * 2048, "ingest_note": "pinned host PCM, submit(h + 1) / collect(h) pipelining, results to pinned host memory every hop",
                        "small_batch": small, "ingest_realtime_streams_per_gpu": int(ns * 64.0 / pct(lat_e2e, 0.99))}

    if rank == 0:
        peak, peak_kind = peaks()
        sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
        k1_bytes = n * 31 * K1_BYTES_PER_FRAME
        achieved = k1_bytes / (k1_ms / 1e3) / 1e9
        k1_traffic, k1_inst, k1_src = k1_ncu()
        K1_WARP_INST_PER_FRAME = k1_inst or 0.0
        line = {"metric": "kws_clips_per_s", "value": value, "unit": "clips/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "q15/int8", "data": "synthetic",
                "config": {"workload": "config#3 shard: %d clips x 31 frames x 1024 int16 per GPU, q15 MFCC + int8 NNoM net" % n,
                           "clip_samples": CLIP, "global_clips": world * n, "l2": "inputs (8.3 GB/GPU) larger than L2",
                           "parallelism": "clip shards, dp%d" % world, "peak": peak_kind, "numa_bound_cpus": numa},
                "mfcc_frames_per_s": world * n * 31 / (k1_ms / 1e3), "audio_seconds_per_s": value * 31 * 0.064,
                # BASELINE.json words the metric in "1 s 16 kHz" clips; the net's own window is 31 frames = 1.984 s
                "one_second_equiv_clips_per_s": value * 31 / 15.625,
                "k1_ms": k1_ms, "k2_ms": k2_ms, "comm_ms": comm_ms, "comm_ms_min_over_ranks": comm_ms_min,
                "comm": "one all_gather of the int8 logits + all_reduce of the int32[10] class counts after the last step (inside the timed region)"
                        if world > 1 else "single GPU: no exchange",
                "parity_checked": parity,
                "cfg2_mfcc": {"workload": "config#2: 65,536 x 1 s clips (15 frames each), q15 MFCC only, 1 GPU",
                              "frames_per_s": cfg2_fps, "clips_per_s": CFG2_CLIPS / (cfg2_ms / 1e3), "ms": cfg2_ms},
                "roofline": {"kernel": "mfcc_q15_kernel", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": k1_traffic, "traffic_source": k1_src, "bytes_per_launch": k1_bytes,
                             # the bound that actually binds: 4 warp-instructions per clock per SM (measured pipe rates, profiles/)
                             "secondary": {"bound": "int-issue", "unit": "G warp-inst/s", "warp_inst_per_frame": K1_WARP_INST_PER_FRAME,
                                           "achieved": n * 31 / (k1_ms / 1e3) * K1_WARP_INST_PER_FRAME / 1e9,
                                           "peak": sm_count * 4 * (clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965) * 1e6 / 1e9,
                                           "frac": (n * 31 / (k1_ms / 1e3) * K1_WARP_INST_PER_FRAME) /
                                                   (sm_count * 4 * (clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965) * 1e6)}},
                "e2e": {"value": e2e_val, "unit": "clips/s", "h2d_bytes_per_step": ne * CLIP * 2, "d2h_bytes_per_step": ne * 20,
                        "clips_per_step": ne, "matches_device_path": same,
                        "bound": "pcie", "h2d_peak_gbs": h2d_gbs, "achieved_gbs": (e2e_val / world) * CLIP * 2 / 1e9,
                        "frac_of_h2d_peak": (e2e_val / world) * CLIP * 2 / 1e9 / h2d_gbs},
                "gpu_launches": int(launches), "clocks": clocks}
        if stream_stats:
            line["streaming"] = stream_stats
        # the CNN's own roofline (BASELINE.json: "% tensor peak"): 784,752 MAC = 1,569,504 int8 ops per clip against the
        # int8 tensor rate MEASURED on this pool (tcgen05.mma.kind::i8 on all SMs, tools/ubench/umma_probe.cu)
        i8_peak, i8_kind = int8_peak()
        k2_tops = n * 1569504 / (k2_ms / 1e3) / 1e12
        line["cnn_roofline"] = {"kernel": "nnom_q7_umma_kernel (tcgen05.mma.kind::i8, TMEM accumulators)", "bound": "tensor", "achieved": k2_tops,
                                "peak": i8_peak, "unit": "TOP/s (int8)", "frac": k2_tops / i8_peak, "peak_kind": i8_kind,
                                "share_of_step": k2_ms / ms_per_step,
                                "note": "algorithmic ops; the kernel issues ~1.6x more (wrap-around pixels, Toeplitz conv1) and is bound by "
                                        "epilogue latency, not by the tensor pipe (DESIGN.md 4)"}
        line["cfg4_float_cnn"] = cfg4
        if generic:
            line["generic_executor"] = generic
        line["float_mfcc"] = dict(float_stats, workload="mfcc_utils.mfcc arithmetic on 65,536 x 1 s clips (15 frames each), 1 GPU")
        if world == 1 and not args.no_cpu:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import cpu_baseline
            ncpu = os.cpu_count() or 1
            # B1 = firmware-true build, B2 = the reference's plain-C CMSIS path at -O3 (AVX2) + NNoM local backend (the fair CPU
            # figure), port = the oracle's plain-C restatement.  All cores and one core each; the object's value is the fastest reference build.
            extra = {}
            for which in ("reference", "reference_o3", "port"):
                for procs in (ncpu, 1):
                    r2 = cpu_baseline.run(clips_per_proc=32 if procs > 1 else 48, procs=procs, reps=3, warmup=1, which=which)
                    extra["%s_%s" % (r2["kind"], "all_cores" if procs > 1 else "1_core")] = {
                        "value": r2["clips_per_s"], "unit": "clips/s", "cores": r2["cores"], "sample": r2["sample"]}
            pick = "reference_o3_all_cores" if "reference_o3_all_cores" in extra else ("reference_all_cores" if "reference_all_cores" in extra else "port_all_cores")
            line["cpu_baseline"] = {"value": extra[pick]["value"], "unit": "clips/s", "cores": extra[pick]["cores"],
                                    "kind": "port" if pick.startswith("port") else "reference", "sample": extra[pick]["sample"],
                                    "build": {"reference_o3_all_cores": "B2: the reference's plain-C CMSIS branches at -O3 -march=x86-64-v3 + NNoM local backend at -O3",
                                              "reference_all_cores": "B1 firmware-true: audioprocessing.c -O0 + CMSIS SIMD branches on emulated M4 intrinsics",
                                              "port_all_cores": "oracle/libedison_oracle.so (plain-C restatement)"}[pick]}
            line["cpu_baseline_builds"] = extra
            line["cpu_baseline_port"] = dict(extra.get("port_all_cores", {}), kind="port")
            # config #1's CPU side: the numpy restatement of mfcc_utils.mfcc (pinned to the reference module's own outputs) on one core
            import mfcc_float as mf
            rng = __import__("numpy").random.default_rng(20)
            xs = rng.integers(-8000, 8000, size=20000 * 1024, dtype="int16")
            t0 = time.perf_counter()
            mf.mfcc_host(xs)
            dt = time.perf_counter() - t0
            line["float_mfcc"]["cpu_numpy"] = {"frames_per_s": 20000 / dt, "cores": 1, "kind": "port", "sample": "20,000 frames, oracle/mfcc_float.py::mfcc_host (vectorised numpy; the reference loops frame by frame)"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
