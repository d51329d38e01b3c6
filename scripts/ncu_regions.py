#!/usr/bin/env python3
"""Sum the warp instructions / samples of an ncu source page over source-line regions.
usage: ncu_regions.py report.ncu-rep kernel-regex file:lo-hi[:name] ...   (lines outside every region -> 'other')"""
import csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
regs = []
for a in sys.argv[3:]:
    p = a.split(":")
    lo, hi = p[1].split("-")
    regs.append((p[0], int(lo), int(hi), p[2] if len(p) > 2 else a))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + pat],
                     capture_output=True, text=True).stdout
fname, hdr = None, None
agg = {}
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < 8 or r[0] == "":
        continue
    try:
        ln, smp, ins = int(r[0]), int(r[6] or 0), int(r[7] or 0)
    except ValueError:
        continue
    name = "other:" + fname
    for f, lo, hi, n in regs:
        if fname == f and lo <= ln <= hi:
            name = n
            break
    a = agg.setdefault(name, [0, 0])
    a[0] += smp
    a[1] += ins
ts = sum(a[0] for a in agg.values()) or 1
ti = sum(a[1] for a in agg.values()) or 1
print(f"total samples {ts} warp-instructions {ti}")
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{a[1]:10d} {a[1] / ti:6.3f} instr | {a[0]:6d} {a[0] / ts:6.3f} samples | {k}")
