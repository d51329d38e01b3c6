#!/usr/bin/env python3
"""Per-source-line summary of an ncu report (cuda,sass source page): samples and warp instructions executed.
usage: ncu_lines.py report.ncu-rep [kernel-regex] [N]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
pat = sys.argv[2] if len(sys.argv) > 2 else None
N = int(sys.argv[3]) if len(sys.argv) > 3 else 40
args = ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"]
if pat:
    args += ["--kernel-name", "regex:" + pat]
out = subprocess.run(args, capture_output=True, text=True).stdout
fname = None
rows = []
hdr = None
seen_launch = 0
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < 8:
        continue
    if r[0] == "":
        continue  # sass row
    try:
        rows.append((fname, int(r[0]), r[1].strip()[:90], int(r[6] or 0), int(r[7] or 0)))
    except ValueError:
        pass
agg = {}
for f, ln, src, smp, ins in rows:
    k = (f, ln)
    a = agg.setdefault(k, [src, 0, 0])
    a[1] += smp
    a[2] += ins
ts = sum(a[1] for a in agg.values()) or 1
ti = sum(a[2] for a in agg.values()) or 1
print(f"total samples {ts}  total warp-instructions {ti}")
print("--- by samples")
for k, a in sorted(agg.items(), key=lambda x: -x[1][1])[:N]:
    print(f"{a[1]:6d} {a[1]/ts:6.3f} | {a[2]:9d} {a[2]/ti:6.3f} | {k[0]}:{k[1]}  {a[0]}")
print("--- by instructions")
for k, a in sorted(agg.items(), key=lambda x: -x[1][2])[:N]:
    print(f"{a[1]:6d} {a[1]/ts:6.3f} | {a[2]:9d} {a[2]/ti:6.3f} | {k[0]}:{k[1]}  {a[0]}")
