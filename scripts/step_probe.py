#!/usr/bin/env python3
"""k_step launch time of the STEPPING instance (what a decoder in the loop uses) and of the replay, device-generated
inputs (not a bench number):   python scripts/step_probe.py [--B 64 512] [--K 1] [--shape c1|owsm] [--L 12] [--mode step|replay|both]

step mode: L lgasr_step launches queued back to back on one stream (no poll, no host sync between them), CUDA events
around the L launches; the first position (one hypothesis per utterance) is excluded by timing positions 1..L-1."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from llm_guided_asr_b200.engine import FusedStepEngine

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, nargs="+", default=[64, 512])
ap.add_argument("--K", type=int, default=1)
ap.add_argument("--shape", default="c1")
ap.add_argument("--L", type=int, default=12)
ap.add_argument("--mode", default="both")
ap.add_argument("--reps", type=int, default=5)
args = ap.parse_args()

dev = torch.device("cuda:0")
peak, _ = bench.peaks()
shape = dict(T=374, V=5000, beam=10, S=15) if args.shape == "c1" else dict(T=749, V=50000, beam=5, S=7)
T, V, W, S, K, L = shape["T"], shape["V"], shape["beam"], shape["S"], args.K, args.L
for B in args.B:
    g = torch.Generator(device=dev).manual_seed(7)
    logp = torch.log_softmax(3.0 * torch.randn(B, T, V, device=dev, generator=g), dim=-1)
    tabs = [torch.log_softmax(2.0 * torch.randn(L, B * W, V, device=dev, generator=g), dim=-1) for _ in range(K)]
    eos = V - 1
    names = ["decoder", "lm", "lm2", "lm3"][:K]
    wts = [0.7, 0.5, 0.2, 0.1][:K]
    eng = FusedStepEngine(B, T, V, W, S, names, "full", True, 0.3, eos, eos, L, 1, False, device=dev)
    mats = [[t[i] for t in tabs] for i in range(L)]
    eng.set_posteriors(logp, [T] * B)
    a_full = 4 * (T * (W * S + 1 + 4 * W) + K * W * V) + 16 * W
    ab = B * a_full
    out = {}
    for mode in (["step", "replay"] if args.mode == "both" else [args.mode]):
        ts = []
        for rep in range(args.reps):
            eng.begin([L] * B, [0] * B, [eos])
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if mode == "step":
                eng.step(mats[0], [0.0] * K, wts)
                e0.record()
                for i in range(1, L):
                    eng.step(mats[i], [0.0] * K, wts)
                e1.record()
                n = L - 1
            else:
                e0.record()
                eng.replay(mats, [0.0] * K, wts, ld=V)
                e1.record()
                n = L
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3 / n)
        out[mode] = min(ts[1:])
    line = f"{args.shape} B={B:4d} K={K}: "
    for mode, us in out.items():
        line += f"{mode} {us:7.2f} us/launch = {ab / us / 1e3:7.1f} GB/s frac {ab / us / 1e3 / peak:.3f} | "
    print(line + f"algorithmic {ab / 1e6:.1f} MB/launch  {eng.launch_info()}", flush=True)
    eng.close()
    del logp, tabs, mats
    torch.cuda.empty_cache()
