#!/usr/bin/env python
"""Markdown rows from bench.py JSON lines (one file per GPU count): the BASELINE.md section 3 / DESIGN.md section 5 tables.
usage: bench_table.py n1.json [n2.json n4.json n8.json]"""
import json, sys
runs = []
for f in sys.argv[1:]:
    j = json.loads(open(f).read().strip().splitlines()[-1])
    runs.append(j)
runs.sort(key=lambda j: j["n_gpus"])
def wl(j, name):
    if name == "c2":
        return {"value": j["value"], "roofline": j["roofline"], "ms_per_step": j["ms_per_step"], "e2e": j.get("e2e"), "selected_records": j["config"].get("selected_records")}
    return j.get("workloads", {}).get(name)
print("| workload | " + " | ".join("%d GPU: pipeline GB/s (scan kernel GB/s = frac of HBM; ms per 100 M records per GPU)" % j["n_gpus"] for j in runs) + " | e2e GB/s (of H2D bound) |")
print("|---|" + "---|" * (len(runs) + 1))
for name in ("c2", "c3", "c4", "c5v", "c5c"):
    cells, e2e = [], []
    for j in runs:
        w = wl(j, name)
        if not w:
            cells.append("—")
            continue
        r = w["roofline"]
        cells.append("%.0f (%.0f = %.3f; %.2f ms)" % (w["value"], r["achieved"], r["frac"], r["pipeline_ms"]))
        e = w.get("e2e")
        if e:
            e2e.append("N=%d: %.1f%s" % (j["n_gpus"], e["value"], (" (%.2f of %.1f)" % (e["frac"], e["bound_gbs"])) if e.get("bound_gbs") else ""))
    print("| %s | %s | %s |" % (name, " | ".join(cells), "; ".join(e2e) or "—"))
j = runs[0]
c = j.get("cpu_baseline")
if c:
    print("\nCPU reference (same run, %d cores): %.2f GB/s with one process per core, %.3f GB/s one process (-O2), %s GB/s one process (-O0)" %
          (c["cores"], c["value"], c["single_process_value"], ("%.3f" % c["single_process_O0_value"]) if c.get("single_process_O0_value") else "n/a"))
cli = j.get("e2e_cli") or {}
for k, v in cli.items():
    if isinstance(v, dict) and "value" in v:
        print("CLI %s: %.1f GB/s whole process (%.2f s for %.1f GB), %.1f GB/s streaming; main thread: %s; reference %.3f GB/s" %
              (k, v["value"], v["process_seconds"], v["bytes"] / 1e9, v.get("streaming_gbs") or 0, v.get("main_thread"), v["reference_process_gbs"]))
