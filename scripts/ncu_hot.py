#!/usr/bin/env python
"""ncu_hot.py REPORT.ncu-rep [N] — the N SASS instructions with the most warp-stall samples, with their stall mix."""
import csv, io, subprocess, sys
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]; data = [r for r in rows[2:] if len(r) == len(hdr)]
iS = hdr.index("# Samples"); iSrc = hdr.index("Source")
stalls = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[iS] or 0) for r in data)
agg = {h: sum(int(r[i] or 0) for r in data) for i, h in stalls}
print("total samples", tot, {h[6:]: v for h, v in sorted(agg.items(), key=lambda x: -x[1])[:8]})
top = sorted(range(len(data)), key=lambda i: -int(data[i][iS] or 0))[:N]
for i in sorted(top):
    r = data[i]
    mix = sorted(((int(r[j] or 0), h[6:]) for j, h in stalls), reverse=True)[:3]
    print("%5d %5s  %-60s %s" % (i, r[iS], r[iSrc].strip()[:60], " ".join(f"{h}={v}" for v, h in mix if v)))
