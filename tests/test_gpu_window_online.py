"""SURVEY.md 8(f) rank 3 on the GPU: the attention-window path of the CTC prefix scorer (``margin > 0`` with attention
weights, ``ctc_prefix_score.py:56-61,143-148``) against vectors recorded from the reference, and an online driver that
calls ``extend_prob`` / ``extend_state`` between blocks the way ``BatchBeamSearchOnline.extend`` does
(``batch_beam_search_online.py:520-535``), with the unmodified reference ``BatchBeamSearch`` doing the search."""

import numpy as np
import pytest
import torch

from test_oracle_window import Z, run_window

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_cuda_scorer_attention_window_matches_reference():
    from llm_guided_asr_b200 import CTCPrefixScorer
    from llm_guided_asr_b200.testing import synthetic as syn

    T, n = int(Z["T"]), int(Z["n"])
    sc = CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=int(Z["eos"]), margin=int(Z["margin"]))
    sc.batch_init_state(torch.from_numpy(Z["logp"]).to(DEV))
    prefixes = [[int(Z["eos"])] for _ in range(n)]

    def score(c, last, olen, ids, att, f_prev, state):
        y = torch.tensor(prefixes, dtype=torch.int64, device=DEV)
        assert y.shape[1] - 1 == olen and y[:, -1].tolist() == last.tolist()
        scores, bst = sc.batch_score_partial(y, torch.from_numpy(ids).to(DEV), [None] * n if state is None else state, None,
                                             att_w=torch.from_numpy(att).to(DEV))
        s = scores.cpu().numpy()
        return np.take_along_axis(s, ids, 1), s[:, sc.eos], bst[6], bst

    def select(bst, hyp, tok):
        sts = [sc.select_state(bst, int(h), int(t)) for h, t in zip(hyp, tok)]
        r_new = np.stack([st[0][:, :T].t().cpu().numpy() for st in sts], axis=2)  # (T, 2, n) like the reference
        s_new = np.array([float(st[1]) for st in sts], np.float32)
        prefixes[:] = [prefixes[int(h)] + [int(t)] for h, t in zip(hyp, tok)]
        return r_new, s_new, sts

    run_window(score, select)


def _have_ref():
    try:
        import espnet.nets.batch_beam_search  # noqa: F401

        return True
    except Exception:
        return False


@pytest.mark.skipif(not _have_ref(), reason="baseline/_ref not installed (scripts/install_ref.sh)")
def test_online_block_driver_extends_between_blocks():
    """A minimal BatchBeamSearchOnline-style driver: the encoder output arrives in growing blocks; before every block the
    scorers are extended (``extend_prob`` / ``extend_state``, batch_beam_search_online.py:520-535), then the unmodified
    reference ``BatchBeamSearch.search`` advances a fixed number of positions.  The CUDA scorer against the reference's
    own scorer in the same driver."""
    from espnet.nets.batch_beam_search import BatchBeamSearch
    from espnet.nets.beam_search import BeamSearch
    from espnet.nets.scorers.ctc import CTCPrefixScorer as RefCTC

    from llm_guided_asr_b200 import CTCPrefixScorer
    from llm_guided_asr_b200.testing import synthetic as syn

    V, beam = 60, 4
    eos = V - 1
    logp, table, _ = syn.peaky_case(91, 66, V, 9, eos)
    blocks = [24, 46, 66]  # frames available after each block
    per_block = 3  # positions decoded per block

    def drive(ctc_scorer, device):
        scorers = dict(decoder=syn.PrefixTableScorer(table.to(device)), ctc=ctc_scorer)
        bs = BeamSearch(beam_size=beam, weights=dict(decoder=0.6, ctc=0.4), scorers=scorers, sos=eos, eos=eos, vocab_size=V,
                        token_list=None, pre_beam_score_key="full")
        bs.__class__ = BatchBeamSearch
        bs.to(device=device, dtype=torch.float32).eval()
        x = logp.to(device)
        trace, ended, pos = [], [], 0
        with torch.no_grad():
            hyps = bs.init_hyp(x[: blocks[0]])
            for bi, nfr in enumerate(blocks):
                h = x[:nfr]
                if bi > 0:  # BatchBeamSearchOnline.extend
                    for k, d in bs.scorers.items():
                        if hasattr(d, "extend_prob"):
                            d.extend_prob(h)
                        if hasattr(d, "extend_state"):
                            hyps.states[k] = d.extend_state(hyps.states[k])
                for _ in range(per_block):
                    best = bs.search(hyps, h)
                    trace.append((best.yseq.cpu().tolist(), best.score.cpu().numpy().copy()))
                    # hypotheses that reached eos leave the running set (batch_beam_search.py:365-429), as in
                    # process_one_block; no length limit here
                    hyps = bs.post_process(pos, 10 ** 6, 0, 1.0, best, ended)
                    pos += 1
        return trace + [([h.yseq.tolist() for h in ended], np.array([float(h.score) for h in ended], np.float32))]

    ref = drive(RefCTC(ctc=syn.IdentityCTC(), eos=eos), "cpu")
    got = drive(CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos), DEV)
    assert len(ref) == len(got) == len(blocks) * per_block + 1
    for (ry, rs), (gy, gs) in zip(ref, got):
        assert ry == gy
        np.testing.assert_allclose(gs, rs, atol=1e-4)
