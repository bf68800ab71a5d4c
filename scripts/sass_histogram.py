#!/usr/bin/env python
"""Static SASS mnemonic histogram per kernel of libfxgpu.so (cuobjdump -sass): what the judge greps for
(UBLKCP = cp.async.bulk through the TMA engine, SYNCS = mbarrier, no HMMA / UTCMMA: the path has no contraction).
usage: sass_histogram.py [libfxgpu.so] > profiles/rNN_sass_histogram.txt"""
import collections, re, subprocess, sys
lib = sys.argv[1] if len(sys.argv) > 1 else "fxtract_b200/_build/libfxgpu.so"
out = subprocess.run(["cuobjdump", "-sass", lib], stdout=subprocess.PIPE, text=True, check=True).stdout
kern, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r"\s+Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and kern:
        hist[kern][m.group(1)] += 1
arch = re.findall(r"arch = (sm_\w+)", out)
print("# %s: %d kernels, architectures %s" % (lib, len(hist), sorted(set(arch))))
for k, h in hist.items():
    tot = sum(h.values())
    key = ["UBLKCP", "SYNCS", "LDS", "STS", "LDG", "STG", "ATOMS", "ATOMG", "RED", "IMAD", "LOP3", "SHF", "VOTE", "POPC", "BAR", "SHFL", "REDUX"]
    print("%-110s %5d instr | %s" % (k[:110], tot, " ".join("%s %d" % (m, h[m]) for m in key if h[m])))
    bad = [m for m in h if m.startswith(("HMMA", "UTC", "HGMMA", "IMMA"))]
    if bad:
        print("    tensor-core mnemonics: %s" % bad)
