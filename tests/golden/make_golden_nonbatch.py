#!/usr/bin/env python3
"""Goldens of the NON-batch path: the unmodified reference ``BeamSearch`` (``espnet/nets/beam_search.py:316-383``)
driving its ``CTCPrefixScorer.init_state / score_partial / select_state`` over the NumPy ``CTCPrefixScore``
(``espnet/nets/ctc_prefix_score.py:272-359``).  Run in the build container only:

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_nonbatch.py

Recorded per case: inputs, the n-best list, and every ``CTCPrefixScore.__call__`` (prefix, candidate ids, previous
state -> log psi and the new states)."""

import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, "/root/reference")
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True

from espnet.nets.beam_search import BeamSearch  # noqa: E402
from espnet.nets.ctc_prefix_score import CTCPrefixScore  # noqa: E402
from espnet.nets.scorers.ctc import CTCPrefixScorer  # noqa: E402
from espnet.nets.scorers.length_bonus import LengthBonus  # noqa: E402

from llm_guided_asr_b200.testing import synthetic as syn  # noqa: E402

CASES = [
    dict(name="nb_peaky", kind="peaky", T=40, V=50, beam=3, ctc=0.4, lm=0.3, pen=0.1, maxr=0.0, ntok=6),
    dict(name="nb_random", kind="random", T=30, V=24, beam=3, ctc=0.3, lm=0.0, pen=0.0, maxr=-8),
    dict(name="nb_ctc_only", kind="peaky", T=36, V=30, beam=2, ctc=1.0, lm=0.0, pen=0.0, maxr=0.0, ntok=5),
]


def run_case(c, seed):
    V, eos = c["V"], c["V"] - 1
    if c["kind"] == "random":
        logp = syn.random_logp(seed, c["T"], V)
        table = syn.random_table(seed + 100, abs(int(c["maxr"])) + 2, 4, V)
    else:
        logp, table, _ = syn.peaky_case(seed, c["T"], V, c["ntok"], eos)
    table_lm = syn.random_table(seed + 200, table.shape[0], 3, V, scale=1.0)
    scorers = dict(decoder=syn.PrefixTableScorer(table), ctc=CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos),
                   length_bonus=LengthBonus(V), lm=syn.PrefixTableScorer(table_lm))
    weights = dict(decoder=1.0 - c["ctc"], ctc=c["ctc"], lm=c["lm"], length_bonus=c["pen"])
    key = None if c["ctc"] == 1.0 else "full"
    bs = BeamSearch(beam_size=c["beam"], weights=weights, scorers=scorers, sos=eos, eos=eos, vocab_size=V, token_list=None,
                    pre_beam_score_key=key)
    bs.eval()
    calls = []
    orig = CTCPrefixScore.__call__

    def rec(self, y, cs, r_prev):
        log_psi, r = orig(self, y, cs, r_prev)
        if len(calls) < 40:
            calls.append(dict(y=np.asarray(y), cs=np.asarray(cs), r_prev=np.array(r_prev), log_psi=np.array(log_psi), r=np.array(r)))
        return log_psi, r

    CTCPrefixScore.__call__ = rec
    try:
        with torch.no_grad():
            nbest = bs(x=logp.clone(), maxlenratio=c["maxr"], minlenratio=0.0)
    finally:
        CTCPrefixScore.__call__ = orig
    out = dict(spec=np.array(json.dumps(dict(c, seed=seed, eos=eos, key=key))), logp=logp.numpy(), table=table.numpy(),
               table_lm=table_lm.numpy(), n_nbest=np.array(len(nbest)), n_calls=np.array(len(calls)))
    L = max(len(h.yseq) for h in nbest)
    out["nbest_yseq"] = np.stack([np.pad(h.yseq.numpy(), (0, L - len(h.yseq)), constant_values=-1) for h in nbest])
    out["nbest_len"] = np.array([len(h.yseq) for h in nbest])
    out["nbest_score"] = np.array([float(h.score) for h in nbest], dtype=np.float32)
    for i, r in enumerate(calls):
        for k, v in r.items():
            out[f"call{i}_{k}"] = v
    return out


if __name__ == "__main__":
    for idx, c in enumerate(CASES):
        out = run_case(c, 3000 + 13 * idx)
        path = os.path.join(HERE, c["name"] + ".npz")
        np.savez_compressed(path, **out)
        print(f"{c['name']:14s} nbest={int(out['n_nbest'])} calls={int(out['n_calls'])} best={float(out['nbest_score'][0]):.4f} "
              f"size={os.path.getsize(path) / 1024:.0f} KB")
