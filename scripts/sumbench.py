#!/usr/bin/env python
"""sumbench.py FILE.json ... — one line per bench.py JSON line (and per entry of its `configs`)."""
import json, sys


def row(name, d):
    if "error" in d:
        print("%-12s ERROR %s" % (name, d["error"]))
        return
    r = d.get("roofline") or {}
    cb = d.get("cpu_baseline") or {}
    f = lambda v, n=3: None if v is None else round(v, n)
    print("%-12s qps=%11.1f ms=%9.4f e2e=%11.1f (%8.4f ms) %-6s kern=%s step=%s L=%s cpu=%s%s clk=%s" % (
        name, d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"].get("ms_per_step", 0), r.get("bound"),
        f(r.get("frac")), f(r.get("step_frac")), d.get("gpu_launches"), f(cb.get("value"), 1),
        "*" if cb.get("extrapolated") else "", (d.get("clocks") or {}).get("sm_mhz")))


for f in sys.argv[1:]:
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
    except Exception as e:
        print(f, "ERR", e)
        continue
    row(f.split("/")[-1][:12], d)
    for name, sub in (d.get("configs") or {}).items():
        row("  " + name, sub)
