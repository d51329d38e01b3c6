"""CPU: the oracle's streaming extension (extend_prob / extend_state) against vectors recorded from the
reference's CTCPrefixScoreTH (tests/golden/make_golden_streaming.py)."""

import os

import numpy as np

from helpers import SCORE_ATOL
from oracle.ctc_prefix_oracle import CTCPrefixOracle

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "streaming.npz"))


def run_streaming(make, score, select, extend_prob, extend_state):
    """The scenario of the golden, parameterised over the implementation under test; returns what the
    golden recorded."""
    eos = int(G["eos"])
    got = {}
    impl = make(G["logp1"])
    sc0, st0 = score(impl, [[eos]], None, G["ids0"])
    got["sc0"] = sc0[0, G["ids0"][0]]
    states = [select(impl, st0, int(h), int(t)) for h, t in G["surv"]]
    got["r_sel"] = np.stack([s[0] for s in states])
    got["s_sel"] = np.array([s[1] for s in states])
    ys = [[eos, 7], [eos, 20]]
    for stage, lp in ((1, G["logp2"]), (2, G["logp3"])):
        extend_prob(impl, lp)
        states = [extend_state(impl, s) for s in states]
        got[f"r_ext{stage}"] = np.stack([s[0] for s in states])
        ids = G[f"ids{stage}"]
        sc, st = score(impl, ys, states, ids)
        got[f"sc{stage}"] = np.take_along_axis(sc, ids, axis=1)
        got[f"sc{stage}_eos"] = sc[:, eos]
        nxt = [(0, 5), (1, 20)]
        states = [select(impl, st, h, t) for h, t in nxt]
        got[f"r_sel{stage}"] = np.stack([s[0] for s in states])
        got[f"s_sel{stage}"] = np.array([s[1] for s in states])
        ys = [ys[0] + [5], ys[1] + [20]]
    return got


def compare(got):
    for k, v in got.items():
        ref = G[k]
        fin = ref > -1e9
        assert ((np.asarray(v) > -1e9) == fin).all(), k
        np.testing.assert_allclose(np.asarray(v)[fin], ref[fin], atol=SCORE_ATOL, rtol=1e-6, err_msg=k)


def test_oracle_streaming_matches_reference():
    eos = int(G["eos"])

    def make(logp):
        return CTCPrefixOracle(logp, 0, eos)

    def score(o, ys, states, ids):
        last = np.array([y[-1] for y in ys])
        if states is None:
            r_prev, s_prev = o.initial_state()
        else:
            r_prev = np.stack([s[0] for s in states], axis=2)
            s_prev = np.array([s[1] for s in states], dtype=np.float32)
        return o(last, len(ys[0]) - 1, r_prev, s_prev, ids)

    def select(o, st, h, t):
        r, s = CTCPrefixOracle.select_state(st, h, t)
        return np.array(r), float(s)

    def extend_state(o, s):
        return o.extend_state(s[0]), s[1]

    compare(run_streaming(make, score, select, lambda o, lp: o.extend_prob(lp), extend_state))
