#!/usr/bin/env python
"""Summarise an ncu launch list: per-kernel count / total / share, and the sequence of one device pass."""
import collections
import csv
import sys

for f in sys.argv[1:]:
    rows = [r for r in csv.reader(open(f)) if len(r) > 5]
    hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
    H = rows[hdr]; ik = H.index("Kernel Name"); iv = H.index("Metric Value"); iu = H.index("Metric Unit")
    seq = []
    for r in rows[hdr + 1:]:
        v = float(r[iv].replace(",", ""))
        v = v / 1e3 if r[iu] == "ns" else (v * 1e3 if r[iu] == "ms" else v)
        seq.append((r[ik].split("(")[0][-40:], v))
    agg = collections.OrderedDict()
    for k, v in seq:
        a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
    tot = sum(a[1] for k, a in agg.items() if "fma_probe" not in k)
    print(f, "total us (without the peak probe)", round(tot, 1))
    for k, (n, t) in agg.items():
        if "fma_probe" in k: continue
        print(f"  {k:42s} n={n:4d} us={t:10.1f} share={t/tot:.3f}")
    i0 = [i for i, (k, _) in enumerate(seq) if "fk_table" in k or "fk_cull" in k]
    if i0:
        print("  first pass:")
        for k, v in seq[i0[0]:i0[0] + 12]: print(f"    {k:42s} {v:9.1f}")
