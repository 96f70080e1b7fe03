#!/usr/bin/env python
"""Per-phase table of the node-line stage kernel from an ncu report: thread-instructions per cell, FP64 and total.
Phases are cut at SASS addresses: prologue (up to the first barrier), A (to the first instruction whose source line
lies in LinePhaseF / LinePointSetup), F (to the second barrier), N (the rest).
Usage: python tools/ncu_phase_table.py prof.ncu-rep libminicfd_b200.so n_cells [kernel-substring]"""
import collections, csv, io, os, re, subprocess, sys, tempfile

rep, lib, n_cells = sys.argv[1], sys.argv[2], int(sys.argv[3])
kname = sys.argv[4] if len(sys.argv) > 4 else "LineStageKernelILi3ELi1"
src = open(os.path.join(os.path.dirname(os.path.abspath(lib)), "csrc", "stage_lines.cuh")).read().splitlines()


def line_of_def(name):
    return next(i + 1 for i, l in enumerate(src) if re.match(r"MCFD_HD \w+ " + name + r"\(", l))


f_begin = line_of_def("LinePointSetup")
f_end = line_of_def("LineLoadUpdate")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
line_at, active, cur = {}, False, None
for l in dis.splitlines():
    if l.startswith(".text."):
        active = kname in l
    elif active:
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
        m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/", l)
        if m:
            line_at[int(m.group(1), 16)] = cur
rows = list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout)))
s = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
h = rows[s]; ci = {k: i for i, k in enumerate(h)}
body = [r for r in rows[s + 1:] if len(r) == len(h)]
base = int(body[0][0], 16)
FP64 = ("DFMA", "DMUL", "DADD", "DSETP", "MUFU")
phase, bars = "prologue", 0
n_bars = sum("BAR.SYNC" in r[1] for r in body)     # the last barrier opens N, the one before it A (an earlier one:
tot = collections.OrderedDict((p, [0, 0, 0]) for p in ("prologue", "A", "F", "N"))
for r in body:
    off = int(r[0], 16) - base
    t = r[1].split()
    op = (t[1] if t and t[0].startswith("@") and len(t) > 1 else (t[0] if t else "")).split(".")[0]
    where = line_at.get(off)
    if phase == "A" and where and where[0] == "stage_lines.cuh" and f_begin <= where[1] < f_end:
        phase = "F"
    thr = int(r[ci["Thread Instructions Executed"]] or 0)
    tot[phase][0] += thr
    tot[phase][1] += thr if op in FP64 else 0
    tot[phase][2] += int(r[ci["# Samples"]] or 0)
    if "BAR.SYNC" in r[1]:
        bars += 1                                   # ... the inter-cell blocks' wait for their ghost states)
        phase = "N" if bars == n_bars else ("A" if bars == n_bars - 1 else phase)
samples = sum(v[2] for v in tot.values()) or 1
print("| phase | thread-instructions per cell | of which FP64 (DFMA DMUL DADD DSETP MUFU) | share of stall samples |")
print("|---|---|---|---|")
for p, (a, f, smp) in tot.items():
    print(f"| {p} | {a / n_cells:.0f} | {f / n_cells:.0f} | {100 * smp / samples:.1f} % |")
a = sum(v[0] for v in tot.values()); f = sum(v[1] for v in tot.values())
print(f"| all | {a / n_cells:.0f} | {f / n_cells:.0f} | 100 % |")
