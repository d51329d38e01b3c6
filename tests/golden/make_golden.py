#!/usr/bin/env python3
"""Generate golden vectors by running the UNMODIFIED reference (imported from /root/reference).

Run in the build container only (the reference does not travel to the GPU box):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

For every case below the reference ``BatchBeamSearch`` (``espnet/nets/batch_beam_search.py``),
wired as ``espnet2/bin/asr_inference.py:173-181,309-315,352-362,379`` wires it, decodes a
synthetic utterance on the CPU in fp32 with the reference ``CTCPrefixScorer`` /
``CTCPrefixScoreTH`` and ``LengthBonus``; the neural full scorers are replaced by the
table scorers of ``llm_guided_asr_b200.testing.synthetic`` (bit-reproducible ``(n, V)`` matrices).
Recorded per case: inputs, n-best (tokens, score, per-scorer scores), the hypotheses kept
after every ``search`` call, and every ``CTCPrefixScoreTH.__call__`` (inputs + outputs).
"""

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

from espnet.nets.batch_beam_search import BatchBeamSearch  # noqa: E402
from espnet.nets.beam_search import BeamSearch  # noqa: E402
from espnet.nets.ctc_prefix_score import CTCPrefixScoreTH  # noqa: E402
from espnet.nets.scorers.ctc import CTCPrefixScorer  # noqa: E402
from espnet.nets.scorers.length_bonus import LengthBonus  # noqa: E402

from llm_guided_asr_b200.testing import synthetic as syn  # noqa: E402

assert syn.BatchScorerInterface.__module__.startswith("espnet."), "reference interfaces not picked up"

CASES = [
    # name, kind, T, V, beam, ctc_w, lm_w, penalty, maxlenratio, minlenratio, extra
    dict(name="an4_shape", kind="random", T=124, V=30, beam=4, ctc=0.3, lm=0.0, pen=0.0, maxr=0.0, minr=0.0),
    dict(name="an4_fixedlen", kind="random", T=124, V=30, beam=4, ctc=0.3, lm=0.0, pen=0.0, maxr=-12, minr=0.0),
    dict(name="lm_bonus", kind="random", T=60, V=200, beam=10, ctc=0.3, lm=0.5, pen=0.1, maxr=-20, minr=0.0),
    dict(name="peaky_end_detect", kind="peaky", T=90, V=120, beam=5, ctc=0.3, lm=0.0, pen=0.0, maxr=0.0, minr=0.0, ntok=12),
    dict(name="peaky_lm", kind="peaky", T=150, V=300, beam=10, ctc=0.3, lm=0.5, pen=0.0, maxr=0.0, minr=0.0, ntok=20),
    dict(name="peaky_minlen", kind="peaky", T=80, V=64, beam=4, ctc=0.5, lm=0.0, pen=0.0, maxr=0.0, minr=0.3, ntok=8),
    dict(name="beam1", kind="peaky", T=70, V=50, beam=1, ctc=0.3, lm=0.0, pen=0.0, maxr=0.0, minr=0.0, ntok=9),
    dict(name="beam2", kind="random", T=40, V=40, beam=2, ctc=0.3, lm=0.0, pen=0.0, maxr=-10, minr=0.0),
    dict(name="no_ctc", kind="peaky", T=60, V=80, beam=5, ctc=0.0, lm=0.3, pen=0.0, maxr=0.0, minr=0.0, ntok=10),
    dict(name="ctc_only", kind="peaky", T=64, V=40, beam=4, ctc=1.0, lm=0.0, pen=0.0, maxr=0.0, minr=0.0, ntok=8),
    dict(name="ctc_only_random", kind="random", T=30, V=24, beam=3, ctc=1.0, lm=0.0, pen=0.0, maxr=-8, minr=0.0),
    dict(name="key_decoder", kind="peaky", T=72, V=90, beam=6, ctc=0.4, lm=0.6, pen=0.2, maxr=0.0, minr=0.0, ntok=10, key="decoder"),
    dict(name="normalize_length", kind="peaky", T=72, V=90, beam=6, ctc=0.3, lm=0.0, pen=0.0, maxr=0.0, minr=0.0, ntok=10, norm=True),
    dict(name="primer", kind="peaky", T=72, V=90, beam=4, ctc=0.3, lm=0.0, pen=0.0, maxr=0.0, minr=0.0, ntok=10, primer=[88, 5, 7]),
    dict(name="tiny_T", kind="random", T=3, V=20, beam=3, ctc=0.3, lm=0.0, pen=0.0, maxr=0.0, minr=0.0),
    dict(name="mid_vocab", kind="random", T=96, V=1000, beam=10, ctc=0.3, lm=0.5, pen=0.0, maxr=-16, minr=0.0),
    # round 2: no CTC scorer + pre_beam_score_key="decoder" + LM: do_pre_beam is False without a partial scorer
    # (beam_search.py:115-119), so the key must be ignored and the top-k taken over the full weighted sum
    dict(name="no_ctc_key_decoder", kind="peaky", T=60, V=80, beam=5, ctc=0.0, lm=0.5, pen=0.1, maxr=0.0, minr=0.0, ntok=10, key="decoder"),
    dict(name="no_ctc_key_random", kind="random", T=40, V=120, beam=6, ctc=0.0, lm=0.7, pen=0.0, maxr=-14, minr=0.0, key="decoder"),
]


def build_inputs(c, seed):
    T, V = c["T"], c["V"]
    eos = V - 1
    if c["kind"] == "random":
        logp = syn.random_logp(seed, T, V)
        P = (abs(int(c["maxr"])) if c["maxr"] < 0 else T) + 2
        table = syn.random_table(seed + 100, min(P, 40), 4, V)
    else:
        logp, table, _ = syn.peaky_case(seed, T, V, c["ntok"], eos)
    table_lm = syn.random_table(seed + 200, table.shape[0], 3, V, scale=1.0)
    return logp, table, table_lm


def run_case(c, seed):
    V, eos = c["V"], c["V"] - 1
    sos = eos  # espnet2 models use sos == eos == V-1
    logp, table, table_lm = build_inputs(c, seed)
    ctc_w = c["ctc"]
    scorers = dict(
        decoder=syn.PrefixTableScorer(table),
        ctc=CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos),
        length_bonus=LengthBonus(V),
        lm=syn.PrefixTableScorer(table_lm),
    )
    weights = dict(decoder=1.0 - ctc_w, ctc=ctc_w, lm=c["lm"], length_bonus=c["pen"])
    key = c.get("key", None if ctc_w == 1.0 else "full")  # asr_inference.py:361
    bs = BeamSearch(
        beam_size=c["beam"],
        weights=weights,
        scorers=scorers,
        sos=sos,
        eos=eos,
        vocab_size=V,
        token_list=None,
        pre_beam_score_key=key,
        normalize_length=c.get("norm", False),
        hyp_primer=c.get("primer"),
    )
    bs.__class__ = BatchBeamSearch  # asr_inference.py:379
    bs.eval()

    calls, steps, kept = [], [], []
    orig_call = CTCPrefixScoreTH.__call__
    orig_search = BatchBeamSearch.search

    def rec_call(self, y, state, scoring_ids=None, att_w=None):
        out, new_state = orig_call(self, y, state, scoring_ids, att_w)
        n = len(y)
        if state is None:
            r_prev = None
            s_prev = np.zeros((n,), np.float32)
        else:
            r_prev = state[0].numpy().copy()
            s_prev = state[1][:, 0].numpy().copy()
        ids = None if scoring_ids is None else scoring_ids.numpy().copy()
        r_new = new_state[0].numpy()
        rec = dict(
            last=np.array([int(yi[-1]) for yi in y]),
            olen=len(y[0]) - 1,
            s_prev=s_prev,
            ids=ids,
            out_eos=out[:, self.eos].numpy().copy(),
        )
        if r_prev is not None:
            rec["r_prev"] = r_prev
        if ids is not None:
            rec["out_ids"] = torch.gather(out, 1, scoring_ids).numpy().copy()
            rec["r_new"] = r_new.copy()  # (T, 2, n, S)
        if out.shape[1] <= 64:
            rec["out_full"] = out.numpy().copy()
        # keep the files small: full traces for the first calls, then every 8th call
        if len(calls) < 6 or len(calls) % 8 == 0:
            rec["index"] = len(calls)
            kept.append(rec)
        calls.append(None)
        return out, new_state

    def rec_search(self, running_hyps, x, pre_x=None):
        best = orig_search(self, running_hyps, x, pre_x)
        steps.append(
            dict(
                yseq=best.yseq.clone().numpy(),
                score=best.score.clone().numpy(),
                scores={k: v.clone().numpy() for k, v in best.scores.items()},
            )
        )
        return best

    CTCPrefixScoreTH.__call__ = rec_call
    BatchBeamSearch.search = rec_search
    try:
        with torch.no_grad():
            nbest = bs(x=logp.clone(), maxlenratio=c["maxr"], minlenratio=c["minr"])
    finally:
        CTCPrefixScoreTH.__call__ = orig_call
        BatchBeamSearch.search = orig_search

    out = dict(
        spec=np.array(json.dumps(dict(c, seed=seed, sos=sos, eos=eos, key=key, names=list(bs.scorers)))),
        logp=logp.numpy(),
        table=table.numpy(),
        table_lm=table_lm.numpy(),
        n_nbest=np.array(len(nbest)),
        n_steps=np.array(len(steps)),
        n_calls=np.array(len(calls)),
        kept_calls=np.array([r["index"] for r in kept]),
    )
    L = max(len(h.yseq) for h in nbest)
    out["nbest_yseq"] = np.stack([np.pad(h.yseq.numpy(), (0, L - len(h.yseq)), constant_values=-1) for h in nbest])
    out["nbest_len"] = np.array([len(h.yseq) for h in nbest])
    out["nbest_score"] = np.array([float(h.score) for h in nbest], dtype=np.float32)
    for k in bs.scorers:
        out[f"nbest_scores_{k}"] = np.array([float(h.scores[k]) for h in nbest], dtype=np.float32)
    for i, s in enumerate(steps):
        out[f"step{i}_yseq"] = s["yseq"]
        out[f"step{i}_score"] = s["score"]
        for k, v in s["scores"].items():
            out[f"step{i}_scores_{k}"] = v
    for r in kept:
        for k, v in r.items():
            if k != "index":
                out[f"call{r['index']}_{k}"] = np.asarray(v)
    return out


def main():
    only = set(sys.argv[1:])  # optional: regenerate only the named cases
    for idx, c in enumerate(CASES):
        if only and c["name"] not in only:
            continue
        seed = 1000 + 17 * idx
        out = run_case(c, seed)
        path = os.path.join(HERE, c["name"] + ".npz")
        np.savez_compressed(path, **out)
        print(
            f"{c['name']:20s} steps={int(out['n_steps'])} nbest={int(out['n_nbest'])} "
            f"best_len={int(out['nbest_len'][0])} best_score={float(out['nbest_score'][0]):.4f} "
            f"size={os.path.getsize(path) / 1024:.0f} KB"
        )


if __name__ == "__main__":
    main()
