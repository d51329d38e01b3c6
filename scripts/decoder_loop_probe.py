"""Whole decode loop with a real attention decoder between the positions (not the replay): the reference's
scorer protocol (per-hypothesis Python state lists, select_state per survivor, one decoder call per utterance)
against the device protocol (DeviceCachedDecoderScorer: one decoder call per position for all B*beam row
slots, lgasr_log_softmax_rows, lgasr_reorder_rows).  Wall clock around decode_batch, CUDA-synchronised,
best of 3 after a warm-up; same search, same decoder weights; the n-best lists must agree."""

import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_guided_asr_b200 import BatchBeamSearch, CTCPrefixScorer, LengthBonus  # noqa: E402
from llm_guided_asr_b200.testing import synthetic as syn  # noqa: E402
from llm_guided_asr_b200.device_cache import DeviceCachedDecoderScorer  # noqa: E402

DEV = torch.device("cuda:0")


def main(B=8, T=374, V=5000, D=256, layers=6, beam=10, L=32):
    eos = V - 1
    g = torch.Generator().manual_seed(0)
    xs = [torch.randn(T, D, generator=g).to(DEV) for _ in range(B)]
    dec = syn.TinyDecoder(V, D=D, layers=layers, heads=4, ff=4 * D, seed=1).eval().to(DEV)

    class ProjCTC(torch.nn.Module):
        def __init__(self):
            super().__init__()
            torch.manual_seed(2)
            self.lo = torch.nn.Linear(D, V)

        def log_softmax(self, hs):
            return torch.log_softmax(self.lo(hs), dim=-1)

    ctc = ProjCTC().eval().to(DEV)
    weights = dict(decoder=0.7, ctc=0.3, length_bonus=0.1)

    def search(scorer):
        return BatchBeamSearch(device_caches=False, scorers=dict(decoder=scorer, ctc=CTCPrefixScorer(ctc=ctc, eos=eos), length_bonus=LengthBonus(V)),
                               weights=weights, beam_size=beam, vocab_size=V, sos=eos, eos=eos, pre_beam_score_key="full")

    out = {}
    for name, sc in (("host-state protocol", dec), ("device caches", DeviceCachedDecoderScorer(dec)),
                     ("device caches + graphs", DeviceCachedDecoderScorer(dec, cuda_graphs=True))):
        bs = search(sc)
        best = 1e9
        for it in range(4):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            res = bs.decode_batch(xs, -L, 0.0)
            torch.cuda.synchronize()
            if it:
                best = min(best, time.perf_counter() - t0)
        out[name] = res
        print(f"{name:24s}: {best * 1e3:8.1f} ms per batch of {B} utterances x {bs.last_steps} positions "
              f"({best / bs.last_steps * 1e6:7.0f} us per position, {B / best:7.1f} utts/s)")
    a = out["host-state protocol"]
    for nm in ("device caches", "device caches + graphs"):
        b = out[nm]
        same = all(x.yseq.tolist() == y.yseq.tolist() for u, v in zip(a, b) for x, y in zip(u, v))
        dmax = max(abs(float(x.score) - float(y.score)) for u, v in zip(a, b) for x, y in zip(u, v))
        print(f"{nm}: n-best lists equal to the host-state protocol: {same}; max |score difference| {dmax:.2e}")


if __name__ == "__main__":
    main()
