#!/usr/bin/env python
"""sumbench.py FILE.json ... — one line per bench.py JSON line."""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.load(open(f))
    except Exception as e:
        print(f, "ERR", e); continue
    r = d.get("roofline") or {}
    pm = d.get("profile_ms") or {}
    ks = {k: round(v, 3) for k, v in pm.items() if v}
    ach, fr = r.get("achieved"), r.get("frac")
    print("%-28s qps=%10.1f ms=%9.4f e2e=%10.1f %s ach=%s frac=%s L=%s %s" % (
        f.split("/")[-1], d["value"], d["ms_per_step"], d["e2e"]["value"], r.get("bound"),
        None if ach is None else round(ach, 1), None if fr is None else round(fr, 3), d.get("gpu_launches"), ks))
