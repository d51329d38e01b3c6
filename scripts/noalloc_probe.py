#!/usr/bin/env python3
"""lgasr_row_logsumexp_noalloc against lgasr_row_logsumexp: time per call (CUDA events, 20 calls) and GB/s of logits read."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from llm_guided_asr_b200 import _lib

lib = _lib.load()
dev = torch.device("cuda:0")
for rows, V, dt in ((640, 5000, torch.float32), (160, 50000, torch.float32), (640, 128256, torch.bfloat16), (4096, 32000, torch.bfloat16)):
    x = (3.0 * torch.randn(rows, V, device=dev)).to(dt)
    out = torch.empty(rows, device=dev)
    st = C.c_void_p(int(torch.cuda.current_stream().cuda_stream))
    line = f"rows {rows:5d} V {V:6d} {str(dt)[6:]:8s}"
    for name in ("lgasr_row_logsumexp", "lgasr_row_logsumexp_noalloc"):
        f = getattr(lib, name)
        call = lambda: _lib.check(f(C.c_void_p(x.data_ptr()), 1 if dt == torch.bfloat16 else 0, C.c_int64(V), C.c_void_p(out.data_ptr()), rows, V, st))
        for _ in range(3):
            call()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            call()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 20
        line += f" | {name[6:]:24s} {us:8.1f} us {x.numel() * x.element_size() / us / 1e3:7.0f} GB/s"
    print(line, flush=True)
