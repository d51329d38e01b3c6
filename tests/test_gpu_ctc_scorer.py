"""GPU parity of the stand-alone CTC prefix scorer kernels (C ABI lgasr_ctc_prefix_score /
lgasr_ctc_select_state) against every CTCPrefixScoreTH.__call__ recorded from the reference."""

import numpy as np
import pytest
import torch

from helpers import GOLDEN_NAMES, SCORE_ATOL, Golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_scorer_calls_match_reference(name):
    from llm_guided_asr_b200 import CTCPrefixScorer
    from llm_guided_asr_b200.testing import synthetic as syn

    g = Golden(name)
    c = g.spec
    if c["ctc"] == 0.0:
        pytest.skip("no CTC scorer in this case")
    dev = torch.device("cuda:0")
    sc = CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=c["eos"])
    sc.batch_init_state(torch.from_numpy(g["logp"]).to(dev))
    T, Tp = sc.impl["T"], sc.impl["Tp"]
    for i in g["kept_calls"]:
        p = f"call{int(i)}_"
        last = torch.from_numpy(g[p + "last"]).to(dev).to(torch.int32)
        olen = int(g[p + "olen"])
        n = last.shape[0]
        s_prev = torch.from_numpy(g[p + "s_prev"]).to(dev)
        ids = torch.from_numpy(g[p + "ids"]).to(dev) if (p + "out_ids") in g else None
        if (p + "r_prev") in g:
            rp = torch.from_numpy(g[p + "r_prev"]).to(dev)  # (T, 2, n)
            r_prev = torch.full((n, 2, Tp), -1e10, device=dev)
            r_prev[:, :, :T] = rp.permute(2, 1, 0)
        else:
            r_prev = sc.impl["r0"].unsqueeze(0).expand(n, 2, Tp).contiguous()
        scores, log_psi = sc._score(last, olen, r_prev.contiguous(), s_prev, ids)
        scores = scores.cpu().numpy()
        np.testing.assert_allclose(scores[:, c["eos"]], g[p + "out_eos"], atol=SCORE_ATOL)
        if (p + "out_full") in g:
            np.testing.assert_allclose(scores, g[p + "out_full"], atol=SCORE_ATOL, rtol=1e-6)
        if ids is not None:
            got = np.take_along_axis(scores, g[p + "ids"], axis=1)
            np.testing.assert_allclose(got, g[p + "out_ids"], atol=SCORE_ATOL, rtol=1e-6)
            # survivor states: r[:, :, h, j] of the reference for every candidate of hypothesis 0..n-1
            ref_r = g[p + "r_new"]  # (T, 2, n, S)
            S = ref_r.shape[3]
            hyp = torch.arange(n, device=dev, dtype=torch.int32).repeat_interleave(S)
            tok = ids.reshape(-1).to(torch.int32)
            r_new = sc._select(r_prev.contiguous(), last, hyp.contiguous(), tok.contiguous(), olen)
            r_new = r_new.cpu().numpy().reshape(n, S, 2, Tp)[:, :, :, :T].transpose(3, 2, 0, 1)
            start = max(olen, 1)
            fin = ref_r[start - 1 :] > -1e9
            np.testing.assert_allclose(r_new[start - 1 :][fin], ref_r[start - 1 :][fin], atol=SCORE_ATOL, rtol=1e-6)
            assert ((r_new[start - 1 :] > -1e9) == fin).all()


def test_lazy_survivor_states_equal_materialised_ones():
    """``select_state`` hands out lazy per-hypothesis states; the next ``batch_score_partial`` computes all survivors in one
    launch (hypothesis / token indices given as device scalars, the way the reference search passes them).  Same scores,
    same states as the one-at-a-time path, and a state that was read as a tuple keeps working."""
    from llm_guided_asr_b200 import CTCPrefixScorer
    from llm_guided_asr_b200.testing import synthetic as syn

    dev = torch.device("cuda:0")
    T, V, eos, n = 57, 40, 39, 4
    g = torch.Generator().manual_seed(3)
    logp = torch.log_softmax(2.0 * torch.randn(T, V, generator=g), dim=-1).to(dev)
    sc = CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos)
    sc.batch_init_state(logp)
    y0 = torch.full((1, 1), eos, dtype=torch.int64, device=dev)
    ids0 = torch.tensor([[5, 9, 17, 23, eos]], dtype=torch.int64, device=dev)
    _, bst0 = sc.batch_score_partial(y0, ids0, [None], None)
    toks = [5, 9, 17, 23]
    y1 = torch.tensor([[eos, t] for t in toks], dtype=torch.int64, device=dev)
    ids1 = torch.tensor([[3, 5, 9, 30, eos]] * n, dtype=torch.int64, device=dev)
    hyp_t = torch.zeros(n, dtype=torch.int64, device=dev)
    tok_t = torch.tensor(toks, dtype=torch.int64, device=dev)
    lazy = [sc.select_state(bst0, hyp_t[k], tok_t[k]) for k in range(n)]  # 0-dim device tensors, never read on the host
    eager = [tuple(sc.select_state(bst0, 0, t)) for t in toks]  # tuple(): materialised one at a time
    s_lazy, b_lazy = sc.batch_score_partial(y1, ids1, lazy, None)
    s_eager, b_eager = sc.batch_score_partial(y1, ids1, eager, None)
    assert torch.equal(s_lazy, s_eager)
    assert torch.equal(b_lazy[0][..., :T], b_eager[0][..., :T]) and torch.equal(b_lazy[1], b_eager[1])  # (frames >= T: padding)
    for lz, eg in zip(lazy, eager):  # after the batched call a lazy state reads like the tuple
        assert torch.equal(lz[0][:, :T], eg[0][:, :T]) and float(lz[1]) == float(eg[1]) and lz[2] == eg[2] == T and len(lz) == 4
    # one more position from survivors of different parents, mixed with an already materialised state
    nxt = [sc.select_state(b_lazy, 2, 30), sc.select_state(b_lazy, 0, 3), sc.select_state(b_lazy, 3, 9)]
    _ = nxt[1][0]
    y2 = torch.tensor([[eos, 17, 30], [eos, 5, 3], [eos, 23, 9]], dtype=torch.int64, device=dev)
    s_mixed, _ = sc.batch_score_partial(y2, ids1[:3], nxt, None)
    ref = [tuple(sc.select_state(b_eager, h, t)) for h, t in ((2, 30), (0, 3), (3, 9))]
    s_ref, _ = sc.batch_score_partial(y2, ids1[:3], ref, None)
    assert torch.equal(s_mixed, s_ref)
