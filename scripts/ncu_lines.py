#!/usr/bin/env python
"""Summarise `ncu --page source --csv --print-source cuda,sass` by source line: share of
executed warp instructions and of stall samples.  usage: ncu_lines.py file.csv [topN]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None
data = []
for r in rows:
    if r and r[0] == "Line No":
        hdr = r
        continue
    if not hdr or len(r) < len(hdr) or r[2] != "-":
        continue
    ix = {k: i for i, k in enumerate(hdr)}
    try:
        ln = int(r[0]); inst = float(r[ix["Instructions Executed"]] or 0); samp = float(r[ix["# Samples"]] or 0)
        thr = float(r[ix["Thread Instructions Executed"]] or 0)
    except ValueError:
        continue
    exc = float(r[ix["L1 Wavefronts Shared Excessive"]] or 0)
    data.append((ln, r[1], inst, samp, thr, exc))
ti = sum(d[2] for d in data) or 1; ts = sum(d[3] for d in data) or 1
print("total warp inst %.3g, samples %d" % (ti, ts))
for d in sorted(data, key=lambda d: -d[3])[:top]:
    print("%4d inst%5.1f%% samp%5.1f%% act%5.1f smemexc%9.0f | %s" % (d[0], 100 * d[2] / ti, 100 * d[3] / ts, d[4] / max(d[2], 1), d[5], d[1].strip()[:100]))
