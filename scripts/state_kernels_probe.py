"""Bandwidth of the kernels either side of the step (lgasr_reorder_rows, lgasr_log_softmax_rows) on one GPU:
CUDA events on the launching stream, 5 warm-ups, inputs rotated through a pool larger than L2 so every timed
launch reads from HBM.  Prints achieved GB/s (algorithmic bytes: one read + one write of every element moved)
against the measured HBM peak of MEASURED_PEAKS.json."""

import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_guided_asr_b200.device_cache import log_softmax_rows  # noqa: E402
from llm_guided_asr_b200.engine import FusedStepEngine  # noqa: E402
from llm_guided_asr_b200.testing import synthetic as syn  # noqa: E402

DEV = torch.device("cuda:0")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    PEAK = 6556.8


def timed(fn, n, warm=5):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(warm + i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / n


def lsm(rows, V, dtype):
    es = 2 if dtype == torch.bfloat16 else 4
    pool = max(2, int(400e6 // (rows * V * (es + 4))) + 1)
    xs = [(3 * torch.randn(rows, V, device=DEV)).to(dtype) for _ in range(pool)]
    outs = [torch.empty(rows, V, device=DEV) for _ in range(pool)]
    t = timed(lambda i: log_softmax_rows(xs[i % pool], outs[i % pool]), 50)
    t_ref = timed(lambda i: torch.log_softmax(xs[i % pool].float(), dim=-1, out=None), 50)
    by = rows * V * (es + 4)
    print(f"log_softmax_rows rows={rows} V={V} {str(dtype)[6:]}: {t * 1e6:8.1f} us  {by / t / 1e9:7.0f} GB/s "
          f"({by / t / 1e9 / PEAK:.2f} of HBM peak)   torch.log_softmax(x.float()): {t_ref * 1e6:8.1f} us")


def reorder(B, W, layers, L, D, cap):
    V, T = 64, 16
    eos = V - 1
    eng = FusedStepEngine(B, T, V, W, int(1.5 * W), ["decoder"], "full", True, 0.3, eos, eos, 8, 1, False, device=DEV)
    logp = torch.stack([syn.random_logp(b, T, V) for b in range(min(B, 8))]).to(DEV).repeat((B + 7) // 8, 1, 1)[:B].contiguous()
    eng.set_posteriors(logp, [T] * B)
    eng.begin([8] * B, [0] * B, [eos])
    for i in range(3):
        eng.step([torch.log_softmax(2 * torch.randn(B * W, V, device=DEV), -1)], [0.0], [0.7])
    nrun = eng.poll(with_parents=False)[0]
    npool = max(2, int(400e6 // (layers * B * W * L * D * 4)) + 1)  # sources rotate through > 3x the L2 size
    srcs = [torch.randn(layers, B * W, L, D, device=DEV) for _ in range(npool)]
    src = srcs[0]
    dsts = [torch.empty(layers, B * W, cap, D, device=DEV) for _ in range(2)]
    t = timed(lambda i: eng.reorder_rows(dsts[i & 1], srcs[i % npool], row_elems=L * D), 30)
    by = 2 * layers * int(nrun.sum()) * L * D * 4
    idx = torch.arange(B * W, device=DEV)
    t_ref = timed(lambda i: torch.index_select(srcs[i % npool], 1, idx), 30)
    print(f"reorder_rows B={B} W={W} layers={layers} L={L} D={D}: {t * 1e6:8.1f} us  {by / t / 1e9:7.0f} GB/s "
          f"({by / t / 1e9 / PEAK:.2f} of HBM peak; {int(nrun.sum())} of {B * W} rows running)   "
          f"torch.index_select of all rows: {t_ref * 1e6:8.1f} us")
    eng.close()


if __name__ == "__main__":
    print(f"HBM peak used: {PEAK} GB/s")
    lsm(640, 5000, torch.float32)
    lsm(640, 5000, torch.bfloat16)
    lsm(160, 50000, torch.float32)
    lsm(160, 50000, torch.bfloat16)
    lsm(640, 128256, torch.bfloat16)
    reorder(64, 10, 6, 32, 512, 64)
    reorder(64, 10, 6, 8, 512, 64)
    reorder(32, 5, 18, 32, 1024, 64)
