#!/usr/bin/env python
"""Summarise an .ncu-rep of the stage kernel: headline metrics, stall reasons, hottest SASS lines.
Usage: python tools/ncu_summary.py gpurun_out/prof_X.ncu-rep [out.md]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
out = open(sys.argv[2], "w") if len(sys.argv) > 2 else sys.stdout


def run(page):
    return subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout


rows = list(csv.reader(io.StringIO(run("raw"))))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "gpu__time_duration.sum", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
print(f"# ncu summary of {rep}\n", file=out)
for d in data:
    print("| metric | value | unit |\n|---|---|---|", file=out)
    for w in want:
        if w in hdr:
            i = hdr.index(w)
            print(f"| {w} | {d[i]} | {units[i]} |", file=out)
    st = {}
    for i, h in enumerate(hdr):
        if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued"):
            st[h.replace("smsp__pcsamp_warps_issue_stalled_", "")] = int(float(d[i] or 0))
    tot = sum(st.values()) or 1
    print("\nstall reasons (pc samples): " + ", ".join(f"{k} {100 * v / tot:.1f}%" for k, v in
          sorted(st.items(), key=lambda kv: -kv[1])[:9]) + "\n", file=out)

rows = list(csv.reader(io.StringIO(run("source"))))
secs = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
if secs:
    s = secs[0]
    e = secs[1] if len(secs) > 1 else len(rows)
    h = rows[s]
    body = [r for r in rows[s + 1:e] if len(r) == len(h)]
    ci = {k: i for i, k in enumerate(h)}
    ops = collections.Counter()
    for r in body:
        t = r[1].split()
        if not t:
            continue
        op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
        ops[op.split(".")[0]] += int(r[ci["Instructions Executed"]] or 0)
    tot = sum(ops.values()) or 1
    print(f"SASS instructions: {len(body)}; executed warp-instructions by opcode: " +
          ", ".join(f"{k} {100 * v / tot:.1f}%" for k, v in ops.most_common(14)) + "\n", file=out)
    print("hottest instructions (samples, address, SASS, stalls):\n", file=out)
    for r in sorted(body, key=lambda r: -int(r[ci["# Samples"]] or 0))[:20]:
        st = {k.replace("stall_", ""): int(r[ci[k]]) for k in h
              if k.startswith("stall_") and "Not" not in k and r[ci[k]] not in ("", "0")}
        print(f"- {r[ci['# Samples']]} `{r[0][-5:]}` `{r[1][:64]}` {st}", file=out)
