"""GPU parity of the streaming extension of the drop-in CTCPrefixScorer (extend_prob / extend_state through
lgasr_ctc_extend_state) against vectors recorded from the reference (tests/golden/streaming.npz)."""

import numpy as np
import pytest
import torch

from test_oracle_streaming import G, compare, run_streaming

pytestmark = pytest.mark.gpu


def test_cuda_scorer_streaming_matches_reference():
    from llm_guided_asr_b200 import CTCPrefixScorer
    from llm_guided_asr_b200.testing import synthetic as syn

    dev = torch.device("cuda:0")
    eos = int(G["eos"])

    def make(logp):
        sc = CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos)
        sc.batch_init_state(torch.from_numpy(logp).to(dev))
        return sc

    def score(sc, ys, states, ids):
        y = torch.tensor(ys, dtype=torch.int64, device=dev)
        st = [None] * len(ys) if states is None else [s[2] for s in states]
        scores, bst = sc.batch_score_partial(y, torch.from_numpy(ids).to(dev), st, None)
        return scores.cpu().numpy(), bst

    def select(sc, bst, h, t):
        st = sc.select_state(bst, h, t)
        T = sc.impl["T"]
        r = st[0][:, :T].t().cpu().numpy()  # (T, 2) like the reference
        return r, float(st[1]), st

    def extend_state(sc, s):
        st = sc.extend_state([s[2]])[0]
        T = sc.impl["T"]
        return st[0][:, :T].t().cpu().numpy(), float(st[1]), st

    compare(run_streaming(make, score, select, lambda sc, lp: sc.extend_prob(torch.from_numpy(lp).to(dev)), extend_state))
