#!/usr/bin/env python3
"""Split a kernel's SASS (ncu source page) into regions of equal execution count and print, per region,
warp-instructions per unit of work, share of instructions and share of stall samples.
usage: ncu_regions.py <rep> <units (e.g. frames per launch)> [min inst/unit]"""
import csv, subprocess, sys
rep, units = sys.argv[1], float(sys.argv[2])
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 5
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, data = rows[1], []
for r in rows[2:]:   # first kernel of the report only
    if r and r[0] == 'Kernel Name':
        break
    if len(r) == len(hdr):
        data.append(r)
iA, iS, iSrc = hdr.index('Instructions Executed'), hdr.index('# Samples'), hdr.index('Source')
groups, cur = [], None
for k, r in enumerate(data):
    ex, s = int(r[iA]), int(r[iS])
    toks = r[iSrc].split()
    op = toks[1] if toks[0].startswith('@') else toks[0]
    key = ex / units
    if cur and abs(cur['key'] - key) <= 0.02 * max(key, 1):
        cur['n'] += 1; cur['ex'] += ex; cur['s'] += s; cur['ops'][op] = cur['ops'].get(op, 0) + 1
    else:
        cur = {'key': key, 'start': k, 'n': 1, 'ex': ex, 's': s, 'ops': {op: 1}}; groups.append(cur)
tot_s = sum(g['s'] for g in groups); tot_ex = sum(g['ex'] for g in groups)
print('total warp-inst/unit %.1f, static instructions %d, samples %d' % (tot_ex / units, len(data), tot_s))
for g in groups:
    if g['ex'] / units > thr:
        top = sorted(g['ops'].items(), key=lambda x: -x[1])[:8]
        print('idx %4d n=%4d x%5.2f  inst/unit=%7.1f (%4.1f%%) samples=%4.1f%%  %s' % (g['start'], g['n'], g['key'], g['ex'] / units, 100 * g['ex'] / tot_ex, 100 * g['s'] / tot_s, top))
