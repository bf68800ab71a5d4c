#!/usr/bin/env python
"""Print SASS (address order) with executed counts around the first instruction matching
(source line, mnemonic).  usage: ncu_sass.py file.csv LINE MNEMONIC [count]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
line, mn = sys.argv[2], sys.argv[3]
count = int(sys.argv[4]) if len(sys.argv) > 4 else 150
hdr = None; out = []; cur = ("?", "?"); curf = "?"
for r in rows:
    if r and r[0] == "File Path": curf = r[1]; continue
    if r and r[0] == "Line No": hdr = r; ix = {k: i for i, k in enumerate(hdr)}; continue
    if not hdr or len(r) < len(hdr): continue
    if r[2] == "-": cur = (curf.split("/")[-1][:14], r[0]); continue
    try: inst = float(r[ix["Instructions Executed"]] or 0)
    except ValueError: continue
    out.append((r[2], r[3], inst, float(r[ix["Thread Instructions Executed"]] or 0), cur))
out.sort(key=lambda x: x[0])
start = 0
for k, o in enumerate(out):
    if o[4][1] == line and mn in o[1]: start = k; break
for o in out[max(0, start - 5):start + count]:
    print("%s %9.0f act%5.1f %-60s %s:%s" % (o[0][-5:], o[2], o[3] / max(o[2], 1), o[1][:60], o[4][0], o[4][1]))
