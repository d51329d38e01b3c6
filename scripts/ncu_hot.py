#!/usr/bin/env python3
"""Summarise an ncu report here (no GPU): raw metrics of each captured launch and the hottest SASS
windows / stall reasons of one kernel.   usage: ncu_hot.py report.ncu-rep [kernel-regex] [window]"""
import csv, io, subprocess, sys

rep = sys.argv[1]
pat = sys.argv[2] if len(sys.argv) > 2 else None
W = int(sys.argv[3]) if len(sys.argv) > 3 else 40

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h = rows[0]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__grid_size", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "sm__cycles_elapsed.max", "lts__t_bytes.sum", "launch__shared_mem_per_block_dynamic"]
idx = [(w, h.index(w)) for w in want if w in h]
for r in rows[2:]:
    print({w: r[i] for w, i in idx})

args = ["ncu", "-i", rep, "--page", "source", "--csv"]
if pat:
    args += ["--kernel-name", "regex:" + pat]
src = subprocess.run(args, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
si = hdr.index("# Samples")
stalls = [i for i, x in enumerate(hdr) if x.startswith("stall_") and "Not Issued" not in x]
inst = []
for r in rows[2:]:
    if len(r) != len(hdr):
        break
    inst.append(r)
tot = sum(int(r[si]) for r in inst) or 1
agg = {hdr[i]: sum(int(r[i] or 0) for r in inst) for i in stalls}
print("instructions", len(inst), "samples", tot)
print(sorted(agg.items(), key=lambda x: -x[1])[:8])
for s in range(0, len(inst), W):
    ss = sum(int(r[si]) for r in inst[s:s + W])
    if ss / tot < 0.02:
        continue
    ops = {}
    for r in inst[s:s + W]:
        t = r[1].split()
        op = t[1] if t[0].startswith("@") else t[0]
        ops[op] = ops.get(op, 0) + 1
    st = {hdr[i]: sum(int(r[i] or 0) for r in inst[s:s + W]) for i in stalls}
    top = sorted(ops.items(), key=lambda x: -x[1])[:5]
    print(s, ss, f"{ss / tot:.3f}", top, sorted(st.items(), key=lambda x: -x[1])[:3])
