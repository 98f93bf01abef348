#!/usr/bin/env python3
"""Print the handful of ncu metrics DESIGN.md / profiles/ quote, from an .ncu-rep.
usage: ncu_summary.py <rep> [--json out.json --frames N [--kernel substr]]   (--json: the per-launch figures bench.py reads, e.g. profiles/k1_ncu.json)"""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum",
        "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_tensor.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print("==", r[hdr.index("Kernel Name")][:60])
    for k in KEYS:
        if k in hdr:
            print("  %-70s %s %s" % (k, r[hdr.index(k)], units[hdr.index(k)]))
    st = [(float(r[i]), h) for i, h in enumerate(hdr)
          if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio") and r[i]]
    print("  stalls per issue:", ", ".join("%s=%.2f" % (h.split("stalled_")[1].split("_per_")[0], v) for v, h in sorted(st, reverse=True)[:8]))

if "--json" in sys.argv:
    import json
    dst = sys.argv[sys.argv.index("--json") + 1]
    frames = int(sys.argv[sys.argv.index("--frames") + 1])
    want = sys.argv[sys.argv.index("--kernel") + 1] if "--kernel" in sys.argv else "mfcc_q15_kernel"
    pick = [r for r in rows[2:] if want in r[hdr.index("Kernel Name")]]
    r = max(pick, key=lambda r: float(r[hdr.index("gpu__time_duration.sum")]))
    num = lambda k: float(r[hdr.index(k)].replace(",", ""))
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    by = lambda k: num(k) * scale.get(units[hdr.index(k)], 1)
    json.dump({"kernel": want, "frames": frames, "dram_bytes_per_launch": by("dram__bytes_read.sum") + by("dram__bytes_write.sum"),
               "warp_inst_per_frame": num("smsp__inst_executed.sum") / frames, "source": sys.argv[1]}, open(dst, "w"), indent=1)
