#!/usr/bin/env python3
"""k_step launch time on the other BASELINE.json shapes (device-generated inputs, back-to-back replay; not a
bench number): configs[2] (two full-scorer matrices) and configs[3] (OWSM: T = 749, V = 50 000, beam 5, two
matrices).  Algorithmic bytes per SURVEY.md §8(d)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from llm_guided_asr_b200.engine import FusedStepEngine

dev = torch.device("cuda:0")
peak, _ = bench.peaks()
SHAPES = [
    ("configs[2] conformer + LM", dict(B=64, T=374, V=5000, beam=10, S=15, K=2)),
    ("configs[3] OWSM", dict(B=32, T=749, V=50000, beam=5, S=7, K=2)),
    ("configs[3] OWSM", dict(B=128, T=749, V=50000, beam=5, S=7, K=2)),
]
L = 12
for name, c in SHAPES:
    B, T, V, W, S, K = c["B"], c["T"], c["V"], c["beam"], c["S"], c["K"]
    g = torch.Generator(device=dev).manual_seed(7)
    logp = torch.log_softmax(3.0 * torch.randn(B, T, V, device=dev, generator=g), dim=-1)
    tabs = [torch.log_softmax(2.0 * torch.randn(L, B * W, V, device=dev, generator=g), dim=-1) for _ in range(K)]
    eos = V - 1
    names = ["decoder", "lm"][:K]
    eng = FusedStepEngine(B, T, V, W, S, names, "full", True, 0.3, eos, eos, L, 1, False, device=dev)
    mats = [[t[i] for t in tabs] for i in range(L)]
    eng.set_posteriors(logp, [T] * B)
    ts = []
    for rep in range(4):
        eng.begin([L] * B, [0] * B, [eos])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.replay(mats, [0.0] * K, [0.7, 0.5][:K], ld=V)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    us = min(ts[1:]) * 1e3 / L
    # A_step = 4*[T_eff*(W*S + 1 + 4*W) + K*W*V] + 16*W per utterance at full beam (first position: one hypothesis)
    a_full = 4 * (T * (W * S + 1 + 4 * W) + K * W * V) + 16 * W
    a_first = 4 * (T * (S + 1 + 4) + K * V) + 16
    ab = B * (a_first + (L - 1) * a_full) / L
    print(f"{name:26s} B={B:4d} T={T} V={V} beam={W} K={K}: {us:8.2f} us/launch  algorithmic {ab / 1e6:7.1f} MB/launch  "
          f"{ab / us / 1e3:7.1f} GB/s  frac {ab / us / 1e3 / peak:.3f}  {eng.launch_info()}", flush=True)
    eng.close()
    del logp, tabs, mats
    torch.cuda.empty_cache()
