#!/usr/bin/env python3
"""Where one 512-utterance batch of the configs[4] leg spends its time (not a bench number): every phase of
bench.run_configs4's decode_fn timed on the host with a device synchronise after it, then the same phases queued
without synchronising (what the leg does)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llm_guided_asr_b200.engine import FusedStepEngine

dev = torch.device("cuda:0")
B, T, V, W, S, K, L = 512, 374, 5000, 10, 15, 2, 64
g = torch.Generator(device=dev).manual_seed(1)
tabs = [torch.log_softmax(2.0 * torch.randn(L, B * W, V, device=dev, generator=g), dim=-1) for _ in range(K)]
logp = torch.log_softmax(3.0 * torch.randn(B, T, V, device=dev, generator=g), dim=-1)
eos = V - 1
lens = [T - (i % 76) for i in range(B)]
eng = FusedStepEngine(B, T, V, W, S, ["decoder", "lm"], "full", True, 0.3, eos, eos, L, 1, False, device=dev)
mats = [[t[i] for t in tabs] for i in range(L)]


def timed(name, fn, sync=True):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = fn()
    t1 = time.perf_counter()
    if sync:
        torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"  {name:16s} host {1e3 * (t1 - t0):7.3f} ms   host+device {1e3 * (t2 - t0):7.3f} ms", flush=True)
    return r


for rep in range(3):
    print(f"rep {rep}")
    timed("set_posteriors", lambda: eng.set_posteriors(logp, lens))
    timed("begin", lambda: eng.begin([L] * B, [0] * B, [eos]))
    timed("replay", lambda: eng.replay(mats, [0.0] * K, [0.7, 0.5], ld=V))
    res = timed("gather_best", eng.gather_best)
torch.cuda.synchronize()
t0 = time.perf_counter()
eng.set_posteriors(logp, lens)
eng.begin([L] * B, [0] * B, [eos])
eng.replay(mats, [0.0] * K, [0.7, 0.5], ld=V)
t1 = time.perf_counter()
res = eng.gather_best()
t2 = time.perf_counter()
print(f"queued: host {1e3 * (t1 - t0):.3f} ms, to results {1e3 * (t2 - t0):.3f} ms, {len(res)} utterances")
