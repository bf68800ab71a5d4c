#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv): launches after the warm-up
steps are grouped by kernel name; prints mean duration per launch and the share of one step.
usage: launch_summary.py launches.csv [launches per step to keep = last N]"""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
seq = []
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    u = r[ui]
    v = v / 1e3 if u in ("ns", "nsecond") else (v if u in ("us", "usecond") else v * 1e3 if u in ("ms", "msecond") else v)
    seq.append((r[ki], v))
# a step starts at every scan_tiles launch that is not a REDO one (template argument false)
starts = [i for i, (k, _) in enumerate(seq) if "scan_tiles_kernel" in k and ("false" in k or "0>" in k) and "synth" not in k]
if not starts:
    starts = [0]
last = seq[starts[-1]:]
tot = sum(v for _, v in last)
print("last step: %d launches, %.1f us" % (len(last), tot))
for k, v in last:
    print("  %9.1f us  %5.1f %%  %s" % (v, 100 * v / tot, k[:110]))
