#!/usr/bin/env python3
"""Which kernels the decoder call of the decode_real leg spends its time in (torch profiler, CUDA activities, eager - no
graphs -, a few positions in the middle of a decode).  Not a bench number."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

import bench
from baseline import ref_harness as rh
from llm_guided_asr_b200 import BatchBeamSearch, CTCPrefixScorer
from llm_guided_asr_b200.device_cache import DeviceCachedDecoderScorer

dev = torch.device("cuda:0")
cfg = dict(bench.CFG)
B, T, V, W, L, D = cfg["B"], cfg["T"], cfg["V"], cfg["beam"], int(os.environ.get("L", "24")), 512
eos = V - 1
g = torch.Generator().manual_seed(0)
xs = [torch.randn(T, D, generator=g).to(dev) for _ in range(B)]
rh.activate()
from espnet2.asr.decoder.transformer_decoder import TransformerDecoder

torch.manual_seed(0)
dec = TransformerDecoder(V, D, attention_heads=8, linear_units=2048, num_blocks=6, dropout_rate=0.0, positional_dropout_rate=0.0,
                         self_attention_dropout_rate=0.0, src_attention_dropout_rate=0.0).eval().to(dev)


class ProjCTC(torch.nn.Module):
    def __init__(self):
        super().__init__()
        torch.manual_seed(2)
        self.lo = torch.nn.Linear(D, V)

    def log_softmax(self, hs):
        return torch.log_softmax(self.lo(hs), dim=-1)


ctc = ProjCTC().eval().to(dev)
sc = DeviceCachedDecoderScorer(dec, cuda_graphs=False)
bs = BatchBeamSearch(scorers=dict(decoder=sc, ctc=CTCPrefixScorer(ctc=ctc, eos=eos)), weights=dict(decoder=0.7, ctc=0.3),
                     beam_size=W, vocab_size=V, sos=eos, eos=eos, pre_beam_score_key="full")
with torch.no_grad():
    bs.decode_batch(xs, -float(L), 0.0)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        bs.decode_batch(xs, -float(L), 0.0)
        torch.cuda.synchronize()
ev = prof.key_averages()
tot = sum(e.device_time_total for e in ev)
print(f"positions {L}  total device time {tot / 1e3:.1f} ms = {tot / 1e3 / L:.3f} ms per position, {sum(e.count for e in ev) / L:.0f} kernels per position")
for e in sorted(ev, key=lambda e: -e.device_time_total)[:28]:
    print(f"{e.device_time_total / tot:6.3f} {e.count / L:7.1f}/pos {e.device_time_total / max(e.count, 1):8.1f} us  {e.key[:110]}")
