"""CPU: the oracle restatement against the golden vectors recorded from the reference itself
(tests/golden/make_golden.py).  This is what pins the oracle (task statement §3)."""

import numpy as np
import pytest

from helpers import GOLDEN_NAMES, SCORE_ATOL, Golden, oracle_from_golden
from oracle.ctc_prefix_oracle import CTCPrefixOracle


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_oracle_nbest_matches_reference(name):
    g = Golden(name)
    c = g.spec
    orc = oracle_from_golden(g)
    orc.trace = []
    nbest = orc.forward(g["logp"], c["maxr"], c["minr"])
    ref = g.nbest()
    assert orc.steps_run == int(g["n_steps"])
    assert len(nbest) == len(ref)
    for h, (yseq, score) in zip(nbest, ref):
        assert h.yseq == yseq  # tokens bit-exact
        assert abs(float(h.score) - score) <= SCORE_ATOL
    for k in c["names"]:
        got = np.array([float(h.scores[k]) for h in nbest], dtype=np.float32)
        np.testing.assert_allclose(got, g[f"nbest_scores_{k}"], atol=2 * SCORE_ATOL, rtol=1e-6)
    # per-step kept hypotheses (order = descending score) are the reference's
    for i, tr in enumerate(orc.trace):
        want = g[f"step{i}_yseq"]
        ys_new = np.concatenate([tr["ys"][tr["top_hyp"]], tr["top_tok"][:, None]], axis=1)
        np.testing.assert_array_equal(ys_new, want[:, : ys_new.shape[1]])
        np.testing.assert_allclose(tr["top_score"], g[f"step{i}_score"], atol=SCORE_ATOL)


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_oracle_ctc_calls_match_reference(name):
    """Every recorded CTCPrefixScoreTH.__call__ replayed through the oracle scorer."""
    g = Golden(name)
    c = g.spec
    if c["ctc"] == 0.0:
        pytest.skip("no CTC scorer in this case")
    ctc = CTCPrefixOracle(g["logp"], 0, c["eos"])
    for i in g["kept_calls"]:
        p = f"call{int(i)}_"
        last, olen, s_prev = g[p + "last"], int(g[p + "olen"]), g[p + "s_prev"]
        ids = g[p + "ids"] if (p + "out_ids") in g else None
        if (p + "r_prev") in g:
            r_prev = g[p + "r_prev"]
        else:
            r_prev = ctc.initial_state()[0]
        scores, (r, log_psi, idmap) = ctc(last, olen, r_prev, s_prev, ids)
        np.testing.assert_allclose(scores[:, c["eos"]], g[p + "out_eos"], atol=SCORE_ATOL)
        if ids is not None:
            got = np.take_along_axis(scores, ids, axis=1)
            np.testing.assert_allclose(got, g[p + "out_ids"], atol=SCORE_ATOL, rtol=1e-6)
            ref_r = g[p + "r_new"]
            fin = ref_r > -1e9
            np.testing.assert_allclose(r[fin], ref_r[fin], atol=SCORE_ATOL, rtol=1e-6)
            assert ((r > -1e9) == fin).all()
        if (p + "out_full") in g:
            np.testing.assert_allclose(scores, g[p + "out_full"], atol=SCORE_ATOL, rtol=1e-6)
