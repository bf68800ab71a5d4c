#!/usr/bin/env python
"""Instruction / stall-sample share per kernel phase.  usage: ncu_phases.py file.csv name=lo-hi ..."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
ranges = {}
for a in sys.argv[2:]:
    k, v = a.split("=")
    lo, hi = v.split("-")
    ranges[k] = (int(lo), int(hi))
hdr = None; cur = None; agg = {}
for r in rows:
    if r and r[0] == "File Path": cur = r[1]; continue
    if r and r[0] == "Line No": hdr = r; ix = {k: i for i, k in enumerate(hdr)}; continue
    if not hdr or len(r) < len(hdr) or r[2] != "-": continue
    try:
        ln = int(r[0]); inst = float(r[ix["Instructions Executed"]] or 0); samp = float(r[ix["# Samples"]] or 0)
        thr = float(r[ix["Thread Instructions Executed"]] or 0)
    except ValueError:
        continue
    if not cur.endswith(".cu"): key = "hdr:" + cur.split("/")[-1]
    else:
        key = "other"
        for name, (a, b) in ranges.items():
            if a <= ln <= b: key = name; break
    x = agg.setdefault(key, [0, 0, 0]); x[0] += inst; x[1] += samp; x[2] += thr
ti = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
print("total warp inst %.4g" % ti)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print("%-32s inst %5.1f%% (%.3g) samples %5.1f%% act %4.1f" % (k, 100 * v[0] / ti, v[0], 100 * v[1] / ts, v[2] / max(v[0], 1)))
