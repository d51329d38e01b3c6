"""The oracle's attention-window path against vectors recorded from the reference's CTCPrefixScoreTH with margin > 0
(tests/golden/att_window.npz, make_golden_window.py)."""

import os

import numpy as np

from helpers import GOLDEN_DIR
from oracle.ctc_prefix_oracle import CTCPrefixOracle

Z = np.load(os.path.join(GOLDEN_DIR, "att_window.npz"))


def run_window(score_fn, select_fn):
    """score_fn(c, last, olen, ids, att, f_prev, state) -> (scores_at_ids (n,S), scores_eos (n,), (f_min, f_max), ctx);
    select_fn(ctx, hyp (n,), tok (n,)) -> (r_new (T,2,n), s_new (n,), next state).  Yields the comparisons."""
    state, f_prev = None, (0, 1)
    for c in range(int(Z["ncalls"])):
        s_ids, s_eos, f, ctx = score_fn(c, Z[f"c{c}_last"], int(Z[f"c{c}_olen"]), Z[f"c{c}_ids"], Z[f"c{c}_att"], f_prev, state)
        assert list(f) == Z[f"c{c}_f"].tolist()
        np.testing.assert_allclose(s_ids, Z[f"c{c}_scores_ids"], atol=1e-4, rtol=1e-6)
        np.testing.assert_allclose(s_eos, Z[f"c{c}_scores_eos"], atol=1e-4, rtol=1e-6)
        r_new, s_new, state = select_fn(ctx, Z[f"c{c}_sel_hyp"], Z[f"c{c}_sel_tok"])
        np.testing.assert_allclose(r_new, Z[f"c{c}_r_new"], atol=1e-4, rtol=1e-6)
        np.testing.assert_allclose(s_new, Z[f"c{c}_s_new"], atol=1e-4, rtol=1e-6)
        f_prev = f


def test_oracle_window_matches_reference():
    T, n = int(Z["T"]), int(Z["n"])
    orc = CTCPrefixOracle(Z["logp"], 0, int(Z["eos"]), margin=int(Z["margin"]))

    def score(c, last, olen, ids, att, f_prev, state):
        if state is None:
            r0, _ = orc.initial_state()
            r_prev, s_prev = np.repeat(r0, n, axis=2), np.zeros(n, np.float32)
        else:
            r_prev, s_prev = state
        scores, st = orc(last, olen, r_prev, s_prev, ids, att_w=att, f_prev=f_prev)
        _, _, f_min, f_max = orc.last_window
        return np.take_along_axis(scores, ids, 1), scores[:, orc.eos], (f_min, f_max), st

    def select(st, hyp, tok):
        rs, ss = zip(*[CTCPrefixOracle.select_state(st, int(h), int(t)) for h, t in zip(hyp, tok)])
        r_new, s_new = np.stack(rs, axis=2), np.array(ss, np.float32)
        return r_new, s_new, (r_new, s_new)

    run_window(score, select)
