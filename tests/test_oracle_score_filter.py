"""CPU: the oracle's restatement of the reference's ScoreFilter against the goldens recorded from the
reference class itself (tests/golden/make_golden_score_filter.py)."""

import os

import numpy as np

from oracle.score_filter_oracle import score_filter, score_filter_mask

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "score_filter.npz")


def test_oracle_matches_reference_goldens():
    g = np.load(GOLD)
    V, sos, eos, notime, first, last = [int(v) for v in g["spec"]]
    assert len(g["ys"]) >= 50
    for y, n, want in zip(g["ys"], g["lens"], g["masks"]):
        got = score_filter_mask(y[:n], V, sos, eos, notime, first, last)
        assert np.array_equal(got, want), y[:n]


def test_batch_form_is_zero_or_minus_inf():
    g = np.load(GOLD)
    V, sos, eos, notime, first, last = [int(v) for v in g["spec"]]
    idx = [i for i in range(len(g["ys"])) if g["lens"][i] == 6]
    out = score_filter(g["ys"][idx, :6], V, sos, eos, notime, first, last)
    assert out.dtype == np.float32 and np.all((out == 0) | np.isneginf(out))
    assert np.array_equal(np.isneginf(out), g["masks"][idx])
