#!/usr/bin/env python
"""mk_results_table.py profiles/r1_bench_*.json — markdown table of the committed bench lines."""
import json, sys
print("| config | queries/s (HBM-resident) | ms / batch | e2e queries/s (host buffers) | dominant kernel | roofline | frac | CPU port q/s (cores) |")
print("|---|---|---|---|---|---|---|---|")
for f in sys.argv[1:]:
    try:
        d = json.load(open(f))
    except Exception:
        continue
    if d.get("impl") == "reference":
        continue
    r = d.get("roofline") or {}
    cb = d.get("cpu_baseline") or {}
    name = f.split("r1_bench_")[-1].replace(".json", "")
    ach = r.get("achieved")
    print("| %s | %s | %.4g | %s | %s | %s %s of %s %s | %s | %s |" % (
        name, f"{d['value']:,.0f}", d["ms_per_step"], f"{d['e2e']['value']:,.0f}", (r.get("kernel") or "").split(" (")[0],
        "" if ach is None else f"{ach:,.0f}", r.get("unit", ""), r.get("peak", ""), r.get("bound", ""),
        "" if r.get("frac") is None else f"{r['frac']:.2f}",
        f"{cb['value']:,.1f} ({cb['cores']})" if cb else "-"))
