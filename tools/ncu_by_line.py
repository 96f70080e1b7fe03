#!/usr/bin/env python
"""Attribute ncu stall samples / executed instructions of the stage kernel to CUDA source lines.
Usage: python tools/ncu_by_line.py prof.ncu-rep libminicfd_b200.so [kernel-substring]
Joins `ncu --page source` (per-SASS-instruction samples, by offset) with `nvdisasm -g` line info."""
import collections, csv, io, os, re, subprocess, sys, tempfile

rep, lib = sys.argv[1], sys.argv[2]
kname = sys.argv[3] if len(sys.argv) > 3 else "StageKernelILi3ELi1"
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
line_of = {}
active, cur = False, None
for l in dis.splitlines():
    if l.startswith(".text."):
        active = kname in l
    elif active:
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
        m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/", l)
        if m:
            line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"],
                                                  capture_output=True, text=True).stdout)))
s = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
h = rows[s]
ci = {k: i for i, k in enumerate(h)}
body = [r for r in rows[s + 1:] if len(r) == len(h)]
base = int(body[0][0], 16)
samples, execd = collections.Counter(), collections.Counter()
for r in body:
    off = int(r[0], 16) - base
    key = line_of.get(off, ("?", 0))
    samples[key] += int(r[ci["# Samples"]] or 0)
    execd[key] += int(r[ci["Instructions Executed"]] or 0)
ts, te = sum(samples.values()), sum(execd.values())
src = {}
for (f, n) in samples:
    p = os.path.join(os.path.dirname(os.path.abspath(lib)), "csrc", f)
    if f not in src and os.path.exists(p):
        src[f] = open(p).read().splitlines()
print(f"total samples {ts}, executed warp-instructions {te}")
for key, v in samples.most_common(int(os.environ.get("TOPN", "45"))):
    f, n = key
    text = src.get(f, [""] * (n + 1))[n - 1].strip()[:70] if n else ""
    print(f"{100 * v / ts:5.1f}% samples {100 * execd[key] / te:5.1f}% instr  {f}:{n}  {text}")
