#!/usr/bin/env python3
"""Phase time line of k_step2 (LGASR_TRACE=1): clock64 stamps of thread 0 of every CTA, cycles since its row stream
started; medians over the CTAs of the last traced position.  usage: trace_step2.py [B] (TRACE_REPLAY=1: graph + PDL)"""
import os, sys
os.environ["LGASR_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from llm_guided_asr_b200 import _lib
from llm_guided_asr_b200.engine import FusedStepEngine

cfg = dict(bench.CFG)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
T, V, W, L = cfg["T"], cfg["V"], cfg["beam"], 8
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(7)
logp = torch.log_softmax(3.0 * torch.randn(B, T, V, device=dev, generator=g), dim=-1)
tables = torch.log_softmax(2.0 * torch.randn(L, B * W, V, device=dev, generator=g), dim=-1)
eos = V - 1
eng = FusedStepEngine(B, T, V, W, cfg["S"], ["decoder"], "full", True, cfg["ctc_weight"], eos, eos, L, 1, False, device=dev)
eng.set_posteriors(logp, [T] * B)
eng.begin([L] * B, [0] * B, [eos])
if os.environ.get("TRACE_REPLAY") == "1":
    eng.replay([[tables[i]] for i in range(L)], [0.0], [1.0 - cfg["ctc_weight"]], ld=V)
else:
    for i in range(L):
        eng.step([tables[i]], [0.0], [1.0 - cfg["ctc_weight"]])
out = np.zeros(B * W * 32, np.int64)
_lib.check(eng.lib.lgasr_debug_trace(eng._h, out.ctypes.data_as(_lib._PI64), out.size))
tr = out.reshape(B * W, 32)
names = ["rows_start", "topS_done", "ctc_in", "state_done", "cand_in", "records", "exchanged", "select_done"]
t0 = tr[:, 0:1]
rel = (tr[:, :8] - t0).astype(np.float64)
sel = tr[:, 7] > tr[:, 6]
for i, n in ((8, "rows_in_regs"), (9, "threshold"), (13, "own_pushes"), (14, "all_pushes")):
    v = (tr[:, i] - tr[:, 0]).astype(np.float64)
    print(f"  row phase: {n:12s} median={np.median(v):9.0f} p10={np.percentile(v, 10):9.0f} p90={np.percentile(v, 90):9.0f} cycles")
for i, n in enumerate(names):
    v = rel[:, i] if i < 7 else rel[sel, i]
    print(f"{n:12s} n={len(v):5d} median={np.median(v):9.0f} p10={np.percentile(v, 10):9.0f} p90={np.percentile(v, 90):9.0f} max={v.max():9.0f} cycles")
d = lambda a, b_: np.median((tr[sel, a] - tr[sel, b_]).astype(np.float64))
print(f"selecting CTAs n={sel.sum()}: exchanged->loaded {d(12, 6):.0f} loaded->ranked {d(10, 12):.0f} ranked->booked {d(11, 10):.0f} booked->done {d(7, 11):.0f} total {d(7, 6):.0f}")
g_entry, g_wait, g_end = tr[:, 16 + 13], tr[:, 16 + 14], tr[:, 16 + 15]
print(f"globaltimer (us): entry spread {(g_entry.max() - g_entry.min()) / 1e3:.2f}; first entry -> last end {(g_end.max() - g_entry.min()) / 1e3:.2f}; "
      f"median CTA life {np.median(g_end - g_entry) / 1e3:.2f}; entry -> past wait median {np.median(g_wait - g_entry) / 1e3:.2f}; "
      f"last non-selecting end {(g_end[~sel].max() - g_entry.min()) / 1e3:.2f}")
print(eng.launch_info())
