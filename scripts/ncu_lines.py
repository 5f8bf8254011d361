#!/usr/bin/env python
"""Aggregate an ncu report's stall samples / executed instructions per CUDA source line.
usage: ncu_lines.py report.ncu-rep kernel-regex [launch-skip] [top]"""
import csv
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name", "regex:" + rx,
                      "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
H = None; cur = None; agg = {}; fname = None
for r in rows:
    if r and r[0] == "File Name": fname = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": H = r; iS = H.index("# Samples"); iI = H.index("Instructions Executed"); continue
    if not H or len(r) < iI + 1: continue
    if r[0] != "": cur = (fname, int(r[0]), r[1][:105]); continue
    try: s = int(r[iS]); i = int(r[iI])
    except ValueError: continue
    a = agg.setdefault(cur, [0, 0]); a[0] += s; a[1] += i
ts = sum(a[0] for a in agg.values()) or 1; ti = sum(a[1] for a in agg.values()) or 1
print("samples", ts, "warp-instr", ti)
for k, (s, i) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{s/ts:6.3f} {i/ti:6.3f} {k[0]}:{k[1]:4d} {k[2]}")
