#!/usr/bin/env python
"""Summarise `ncu --page source --csv --print-source cuda,sass` per kernel and source line: share of
executed warp instructions and of stall samples.  usage: ncu_lines.py file.csv [min_inst_pct] [kernel substring]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.3
want = sys.argv[3] if len(sys.argv) > 3 else ""
kern = None; hdr = None; cur = None
data = collections.OrderedDict()
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": kern = r[1]; data.setdefault(kern, {}); continue
    if r[0] == "Line No": hdr = r; ix = {k: i for i, k in enumerate(hdr)}; continue
    if not hdr or len(r) < len(hdr) or r[2] != "-": continue
    try:
        ln = int(r[0]); inst = float(r[ix["Instructions Executed"]] or 0); samp = float(r[ix["# Samples"]] or 0)
        thrd = float(r[ix["Thread Instructions Executed"]] or 0)
    except (ValueError, KeyError):
        continue
    key = (cur, ln)
    d = data[kern].setdefault(key, [0, 0, 0, r[1].strip()[:95]])
    d[0] += inst; d[1] += samp; d[2] += thrd
for k, d in data.items():
    if want not in k: continue
    ti = sum(v[0] for v in d.values()) or 1; ts = sum(v[1] for v in d.values()) or 1
    print("== %s: %.4g warp inst, %d samples" % (k, ti, ts))
    for (f, ln), v in sorted(d.items(), key=lambda kv: (kv[0][0] != "fxg_scan.cu", kv[0][0], kv[0][1])):
        if 100 * v[0] / ti >= thr or 100 * v[1] / ts >= 2 * thr:
            print("%-14s %4d inst%5.1f%% samp%5.1f%% act%5.1f | %s" % (f[:14], ln, 100 * v[0] / ti, 100 * v[1] / ts, v[2] / max(v[0], 1), v[3]))
