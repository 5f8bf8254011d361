#!/usr/bin/env python
"""Per source-line RANGE: share of warp instructions, lanes per instruction (predicated-on threads / instructions) and
stall samples, from an ncu report's source page.  usage: ncu_regions.py report.ncu-rep kernel-regex launch-skip lo-hi [lo-hi ...]"""
import csv, subprocess, sys
rep, rx, skip = sys.argv[1], sys.argv[2], sys.argv[3]
ranges = [tuple(int(x) for x in a.split("-")) for a in sys.argv[4:]]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name", "regex:" + rx,
                      "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
H = None; cur = None; agg = {}
for r in csv.reader(out.splitlines()):
    if r and r[0] == "Line No":
        H = r; iS = H.index("# Samples"); iI = H.index("Instructions Executed"); iT = H.index("Predicated-On Thread Instructions Executed"); continue
    if not H or len(r) <= iT: continue
    if r[0] != "": cur = int(r[0]); continue
    try: s, i, t = int(r[iS]), int(r[iI]), int(r[iT])
    except ValueError: continue
    a = agg.setdefault(cur, [0, 0, 0]); a[0] += s; a[1] += i; a[2] += t
ts = sum(a[0] for a in agg.values()) or 1; ti = sum(a[1] for a in agg.values()) or 1; tt = sum(a[2] for a in agg.values())
print(f"all: lanes {tt/ti:.1f}")
for lo, hi in ranges:
    s = sum(a[0] for l, a in agg.items() if lo <= l <= hi); i = sum(a[1] for l, a in agg.items() if lo <= l <= hi); t = sum(a[2] for l, a in agg.items() if lo <= l <= hi)
    print(f"lines {lo}-{hi}: samples {s/ts:.3f} instr {i/ti:.3f} lanes {t/max(i,1):.1f}")
