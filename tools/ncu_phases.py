#!/usr/bin/env python
"""Split the pc samples and executed instructions of a kernel at its barriers (ncu --page source).
Usage: python tools/ncu_phases.py prof.ncu-rep"""
import csv, io, subprocess, sys
rows = list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout)))
s = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
h = rows[s]; ci = {k: i for i, k in enumerate(h)}
body = [r for r in rows[s + 1:] if len(r) == len(h)]
base = int(body[0][0], 16)
stall_cols = [k for k in h if k.startswith("stall_") and "Not" not in k]
seg = []; cur = {"name": "start", "samples": 0, "inst": 0, "n": 0, "stalls": {}}
for r in body:
    off = int(r[0], 16) - base
    cur["samples"] += int(r[ci["# Samples"]] or 0); cur["inst"] += int(r[ci["Instructions Executed"]] or 0); cur["n"] += 1
    for k in stall_cols:
        v = int(r[ci[k]] or 0)
        if v: cur["stalls"][k] = cur["stalls"].get(k, 0) + v
    if "BAR.SYNC" in r[1] or ("BRA" in r[1] and "@" in r[1] and False):
        seg.append(cur); cur = {"name": "after barrier @%x" % off, "samples": 0, "inst": 0, "n": 0, "stalls": {}}
seg.append(cur)
T = sum(x["samples"] for x in seg); I = sum(x["inst"] for x in seg)
print(f"total samples {T}, executed warp-instructions {I}")
for x in seg:
    st = sorted(x["stalls"].items(), key=lambda kv: -kv[1])[:6]
    print(f"{x['name']}: {x['n']} SASS instr, {100 * x['samples'] / T:.1f}% samples, {100 * x['inst'] / I:.1f}% instr, " +
          ", ".join(f"{k.replace('stall_', '')} {v}" for k, v in st))
