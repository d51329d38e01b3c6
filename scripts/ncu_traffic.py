#!/usr/bin/env python3
"""profiles/r02_kstep_traffic.json from an ncu --set full report: DRAM bytes, duration and the counters the judge reads
of the step-kernel launches in the report.   usage: ncu_traffic.py report.ncu-rep kernel-regex out.json [note]"""
import csv, io, json, subprocess, sys
rep, pat, outp = sys.argv[1], sys.argv[2], sys.argv[3]
note = sys.argv[4] if len(sys.argv) > 4 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--kernel-name", "regex:" + pat], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h = rows[0]
units = dict(zip(h, rows[1]))
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "nsecond": 1.0, "ns": 1.0, "usecond": 1e3, "us": 1e3, "msecond": 1e6, "ms": 1e6}
recs = []
for r in rows[2:]:
    d = dict(zip(h, r))
    f = lambda k: float(d[k].replace(",", "")) * SCALE.get(units.get(k, ""), 1.0) if k in d and d[k] not in ("", "n/a") else None
    stalls = {k.split("stalled_")[1]: float(d[k]) for k in h if "pcsamp_warps_issue_stalled" in k and "not_issued" not in k and d[k]}
    tot = sum(stalls.values()) or 1
    recs.append({
        "kernel": d["Kernel Name"], "grid": d.get("launch__grid_size"), "block": d.get("launch__block_size"),
        "duration_us_under_ncu": f("gpu__time_duration.sum") / 1e3 if f("gpu__time_duration.sum") else None,
        "dram_bytes_read": f("dram__bytes_read.sum"), "dram_bytes_write": f("dram__bytes_write.sum"),
        "registers": f("launch__registers_per_thread"), "dyn_smem": f("launch__shared_mem_per_block_dynamic"),
        "warps_active_pct": f("sm__warps_active.avg.pct_of_peak_sustained_active"),
        "ipc_active": f("sm__inst_executed.avg.per_cycle_active"), "issue_active_pct": f("smsp__issue_active.avg.pct"),
        "inst_executed": f("smsp__inst_executed.sum"),
        "stall_share": {k: round(v / tot, 4) for k, v in sorted(stalls.items(), key=lambda x: -x[1])[:8]},
    })
n = len(recs)
out = {
    "source": rep.split("/")[-1], "kernel_regex": pat, "launches": n, "note": note,
    "dram_bytes_per_launch": sum((r["dram_bytes_read"] or 0) + (r["dram_bytes_write"] or 0) for r in recs) / max(n, 1),
    "records": recs,
}
json.dump(out, open(outp, "w"), indent=1)
print(json.dumps({k: out[k] for k in ("launches", "dram_bytes_per_launch")}), recs[0]["stall_share"] if recs else None)
