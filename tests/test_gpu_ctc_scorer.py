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
