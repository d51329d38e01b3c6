#!/usr/bin/env python3
"""k_step launch time vs utterances in flight (device-generated inputs; not a bench number):
   python scripts/scale_probe.py [B ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from llm_guided_asr_b200.engine import FusedStepEngine

Bs = [int(a) for a in sys.argv[1:]] or [64, 128, 256, 512]
cfg = dict(bench.CFG)
T, V, W, L = cfg["T"], cfg["V"], cfg["beam"], 16
dev = torch.device("cuda:0")
peak, _ = bench.peaks()
for B in Bs:
    g = torch.Generator(device=dev).manual_seed(7)
    logp = torch.log_softmax(3.0 * torch.randn(B, T, V, device=dev, generator=g), dim=-1)
    tables = torch.log_softmax(2.0 * torch.randn(L, B * W, V, device=dev, generator=g), dim=-1)
    eos = V - 1
    eng = FusedStepEngine(B, T, V, W, cfg["S"], ["decoder"], "full", True, cfg["ctc_weight"], eos, eos, L, 1, False, device=dev)
    mats = [[tables[i]] for i in range(L)]
    xl, mx, mn = [T] * B, [L] * B, [0] * B
    eng.set_posteriors(logp, xl)
    ts = []
    for rep in range(4):
        eng.begin(mx, mn, [eos])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.replay(mats, [0.0], [1.0 - cfg["ctc_weight"]], ld=V)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    us = min(ts[1:]) * 1e3 / L
    nrun = [[1] * B] + [[W] * B] * (L - 1)
    ab = bench.algorithmic_bytes(dict(cfg, B=B), nrun) / L
    print(f"B={B:4d}  {us:8.2f} us/launch  algorithmic {ab / 1e6:7.1f} MB/launch  {ab / us / 1e3:7.1f} GB/s  frac {ab / us / 1e3 / peak:.3f}  {eng.launch_info()}", flush=True)
    eng.close()
    del logp, tables, mats
    torch.cuda.empty_cache()
