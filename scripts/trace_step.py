#!/usr/bin/env python3
"""Phase time line of k_step on the bench workload (LGASR_TRACE=1): cycles since CTA start, median over CTAs."""
import os, sys
os.environ["LGASR_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import numpy as np, torch
import bench
from llm_guided_asr_b200 import _lib
from llm_guided_asr_b200.engine import FusedStepEngine

cfg = dict(bench.CFG)
if len(sys.argv) > 1:
    cfg["B"] = int(sys.argv[1])
B, T, V, W, L = cfg["B"], cfg["T"], cfg["V"], cfg["beam"], 12
dev = torch.device("cuda:0")
logp, tables = bench.make_inputs(0, dict(cfg, L=L))
d_logp, d_tables = logp.to(dev), tables.to(dev)
eos = V - 1
eng = FusedStepEngine(B, T, V, W, cfg["S"], ["decoder"], "full", True, cfg["ctc_weight"], eos, eos, L, 1, False, device=dev)
eng.set_posteriors(d_logp, [T] * B)
eng.begin([L] * B, [0] * B, [eos])
if os.environ.get("TRACE_REPLAY") == "1":  # graph + PDL launches, as the bench's timed region
    eng.replay([[d_tables[i]] for i in range(L)], [0.0], [1.0 - cfg["ctc_weight"]], ld=V)
else:
    for i in range(L):
        eng.step([d_tables[i]], [0.0], [1.0 - cfg["ctc_weight"]])
out = np.zeros(B * W * 32, np.int64)
_lib.check(eng.lib.lgasr_debug_trace(eng._h, out.ctypes.data_as(_lib._PI64), out.size))
tr = out.reshape(B * W, 2, 16)
t0 = tr[:, 0, 0:1]
names0 = ["start", "ctc_in", "ctc_done", "joined", "cand_in", "records", "ticket", "select_done", "scan1_done", "scan2_done", "sel_ranked", "sel_booked", "sel_loaded"]
names1 = ["start", "row0_in", "sel_done", "thr_done", "chunk0_done"]
rel0 = None  # filled after the selection stamps have been masked
# CTAs that ran the selection in THIS launch (slots 7, 10-12 of the others hold stamps of earlier launches)
last = (tr[:, 0, 7] > tr[:, 0, 6]) & (tr[:, 0, 7] - tr[:, 0, 6] < 1_000_000) & (tr[:, 0, 12] > tr[:, 0, 6])
for i in (7, 10, 11, 12):
    tr[~last, 0, i] = 0
rel0 = (tr[:, 0, :13] - t0).astype(np.float64)
rel1 = (tr[:, 1, :5] - t0).astype(np.float64)
for i, n in enumerate(names0):
    v = rel0[:, i][tr[:, 0, i] != 0]
    if len(v):
        print(f"ctc/all {n:12s} n={len(v):4d} median={np.median(v):9.0f} p90={np.percentile(v, 90):9.0f} max={v.max():9.0f} cycles")
for i, n in enumerate(names1):
    v = rel1[:, i][tr[:, 1, i] != 0]
    if len(v):
        print(f"sel     {n:12s} n={len(v):4d} median={np.median(v):9.0f} p90={np.percentile(v, 90):9.0f} max={v.max():9.0f} cycles")


d = lambda a, b_: np.median((tr[last, 0, a] - tr[last, 0, b_]).astype(np.float64))
print(f"select CTAs: n={last.sum()} ticket->loaded {d(12, 6):.0f}  loaded->ranked {d(10, 12):.0f}  ranked->booked {d(11, 10):.0f}  booked->done {d(7, 11):.0f}  total {d(7, 6):.0f}")

# globaltimer stamps (CTA start / end) alternate between the two rows by step parity: the last two launches
cur, prv = (L - 1) & 1, L & 1
g0, g1 = tr[:, cur, 14], tr[:, cur, 15]
p0, p1 = tr[:, prv, 14], tr[:, prv, 15]
early = np.sort(g0 - p1.max())
print(f"consecutive launches: previous grid first start -> last end {(p1.max() - p0.min()) / 1e3:.2f} us; this grid's first CTA starts "
      f"{(g0.min() - p1.max()) / 1e3:+.2f} us after the previous grid's last CTA ended (median CTA {np.median(early) / 1e3:+.2f} us, last {early[-1] / 1e3:+.2f} us); "
      f"period (last end -> last end) {(g1.max() - p1.max()) / 1e3:.2f} us; previous grid's non-select CTAs all ended {(p1.max() - np.sort(p1)[int(0.9 * len(p1)) - 1]) / 1e3:.2f} us before its last end")
en, pen = tr[:, cur, 13], tr[:, prv, 13]
q = lambda a: " ".join(f"{v / 1e3:+.2f}" for v in np.percentile(a, [0, 10, 50, 90, 100]))
print(f"kernel entry (before griddepcontrol.wait), us relative to the previous grid's last CTA end, min/p10/p50/p90/max: this grid {q(en - p1.max())}; "
      f"previous grid's entries {q(pen - p1.max())}, its post-wait starts {q(p0 - p1.max())}, its ends {q(p1 - p1.max())}")
print(f"this grid: post-wait start {q(g0 - p1.max())}; end {q(g1 - p1.max())}")
# per utterance: the chain from one position's selection to the next one's
G0, G1, EN, P1 = g0.reshape(B, W), g1.reshape(B, W), en.reshape(B, W), p1.reshape(B, W)
u_prev_end, u_end = P1.max(axis=1), G1.max(axis=1)
u_start = G0.max(axis=1)
print(f"per utterance (us, medians): previous selection end -> all CTAs past the hand-off {np.median(u_start - u_prev_end) / 1e3:.2f}; "
      f"hand-off -> selection end {np.median(u_end - u_start) / 1e3:.2f}; period (selection end -> selection end) {np.median(u_end - u_prev_end) / 1e3:.2f} "
      f"(p10 {np.percentile(u_end - u_prev_end, 10) / 1e3:.2f}, p90 {np.percentile(u_end - u_prev_end, 90) / 1e3:.2f}); "
      f"entry -> past the hand-off (selection phase + waiting) {np.median(G0 - EN) / 1e3:.2f}; entry - previous selection end {np.median(EN - u_prev_end[:, None]) / 1e3:+.2f}")
print(f"globaltimer: CTA starts spread {(g0.max() - g0.min()) / 1e3:.2f} us; first start -> last end {(g1.max() - g0.min()) / 1e3:.2f} us; "
      f"median CTA life {np.median(g1 - g0) / 1e3:.2f} us; last ticket-CTA end - first start: {(g1[~last].max() - g0.min()) / 1e3:.2f} us")

jd = tr[:, 0, 3] - np.maximum(tr[:, 0, 2], tr[:, 1, 2])
print(f"join + state write-out after the later of (ctc_done, sel_done): median {np.median(jd):.0f} p90 {np.percentile(jd, 90):.0f} cycles; "
      f"ctc later than sel in {np.mean(tr[:, 0, 2] > tr[:, 1, 2]) * 100:.0f}% of CTAs; cand_in - joined median {np.median(tr[:, 0, 4] - tr[:, 0, 3]):.0f}; "
      f"cand_in - sel_done median {np.median(tr[:, 0, 4] - tr[:, 1, 2]):.0f}")
