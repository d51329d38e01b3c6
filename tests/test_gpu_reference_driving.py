"""INTEGRATION.md level 1 ("scorer only"): the UNMODIFIED reference search objects drive this package's CUDA
``CTCPrefixScorer`` through the plug-in API, and the non-batch scorer API is checked call by call.

* every recorded ``CTCPrefixScore.__call__`` of the reference's NumPy scorer (``ctc_prefix_score.py:272-359``) is
  replayed through ``CTCPrefixScorer.score_partial`` (no reference needed at test time);
* with ``baseline/_ref`` present (``scripts/install_ref.sh``; it travels to the GPU box), the reference's own
  ``BeamSearch`` (non-batch, ``beam_search.py:316-383``) and ``BatchBeamSearch`` (``batch_beam_search.py``) run on
  ``cuda`` with this package's scorer in place of theirs and must reproduce the goldens recorded from the reference.

Bar: tokens bit-exact, scores / log psi / states within 1e-4 abs (fp32; log-zero entries relative 1e-6).
"""

import json
import os

import numpy as np
import pytest
import torch

from helpers import GOLDEN_DIR, NONBATCH_NAMES, SCORE_ATOL, Golden

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
LOGZERO = -1.0e10


def _have_ref():
    try:
        import espnet.nets.batch_beam_search  # noqa: F401

        return True
    except Exception:
        return False


def _ours_from_ref_state(r_prev, Tp):
    """reference state (T, 2) = (r^n, r^b) -> the stand-alone scorer's (2, Tp) rows (r^n, r^b), frames contiguous."""
    T = r_prev.shape[0]
    out = np.full((2, Tp), LOGZERO, np.float32)
    out[0, :T] = r_prev[:, 0]
    out[1, :T] = r_prev[:, 1]
    return out


@pytest.mark.parametrize("name", NONBATCH_NAMES)
def test_score_partial_replays_reference_numpy_scorer_calls(name):
    from llm_guided_asr_b200 import CTCPrefixScorer
    from llm_guided_asr_b200.testing import synthetic as syn

    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    spec = json.loads(str(z["spec"]))
    x = torch.from_numpy(z["logp"]).to(DEV)
    sc = CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=spec["eos"])
    prev0, r0 = sc.init_state(x)
    T, Tp = sc.impl["T"], sc.impl["Tp"]
    assert float(prev0) == 0.0
    for i in range(int(z["n_calls"])):
        y, cs, r_prev = z[f"call{i}_y"], z[f"call{i}_cs"], z[f"call{i}_r_prev"]
        state = (torch.zeros((), device=DEV), torch.from_numpy(_ours_from_ref_state(r_prev, Tp)).to(DEV))
        tscore, (psi, r_new) = sc.score_partial(torch.from_numpy(y).to(DEV), torch.from_numpy(cs).to(DEV), state, x)
        want = z[f"call{i}_log_psi"]
        np.testing.assert_allclose(psi.cpu().numpy(), want, atol=SCORE_ATOL, rtol=1e-6)
        np.testing.assert_allclose(tscore.cpu().numpy(), want, atol=SCORE_ATOL, rtol=1e-6)  # prev_score is 0 here
        r_ref = z[f"call{i}_r"]  # (S, T, 2)
        got = r_new.cpu().numpy()  # (S, 2, Tp)
        special = (cs == spec["eos"]) | (cs == 0)
        # the reference leaves the frames before start - 1 uninitialised (np.ndarray, ctc_prefix_score.py:314-320):
        # only frames >= start - 1 carry state
        lo = max(len(y) - 1, 1) - 1
        for j in range(len(cs)):
            if special[j]:
                continue  # the reference's eos / blank "states" are never selected (the hypothesis ends)
            np.testing.assert_allclose(got[j, 0, lo:T], r_ref[j, lo:, 0], atol=SCORE_ATOL, rtol=1e-6)
            np.testing.assert_allclose(got[j, 1, lo:T], r_ref[j, lo:, 1], atol=SCORE_ATOL, rtol=1e-6)


def _scorers(table, table_lm, V, eos):
    from espnet.nets.scorers.length_bonus import LengthBonus

    from llm_guided_asr_b200 import CTCPrefixScorer
    from llm_guided_asr_b200.testing import synthetic as syn
    from llm_guided_asr_b200.scorer_interface import USING_REFERENCE_INTERFACES

    assert USING_REFERENCE_INTERFACES, "conftest.py puts baseline/_ref on the path before the package is imported"
    return dict(decoder=syn.PrefixTableScorer(torch.from_numpy(table).to(DEV)),
                ctc=CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos), length_bonus=LengthBonus(V),
                lm=syn.PrefixTableScorer(torch.from_numpy(table_lm).to(DEV)))


@pytest.mark.skipif(not _have_ref(), reason="baseline/_ref not installed (scripts/install_ref.sh)")
@pytest.mark.parametrize("name", NONBATCH_NAMES)
def test_reference_nonbatch_beam_search_drives_cuda_scorer(name):
    from espnet.nets.beam_search import BeamSearch

    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    c = json.loads(str(z["spec"]))
    V, eos = c["V"], c["eos"]
    weights = dict(decoder=1.0 - c["ctc"], ctc=c["ctc"], lm=c["lm"], length_bonus=c["pen"])
    bs = BeamSearch(beam_size=c["beam"], weights=weights, scorers=_scorers(z["table"], z["table_lm"], V, eos), sos=eos, eos=eos,
                    vocab_size=V, token_list=None, pre_beam_score_key=c["key"])
    bs.to(device=DEV, dtype=torch.float32).eval()
    with torch.no_grad():
        nbest = bs(x=torch.from_numpy(z["logp"]).to(DEV), maxlenratio=c["maxr"], minlenratio=0.0)
    assert len(nbest) == int(z["n_nbest"])
    for i, h in enumerate(nbest):
        ln = int(z["nbest_len"][i])
        assert h.yseq.tolist() == z["nbest_yseq"][i, :ln].tolist()
        assert abs(float(h.score) - float(z["nbest_score"][i])) <= SCORE_ATOL


@pytest.mark.skipif(not _have_ref(), reason="baseline/_ref not installed (scripts/install_ref.sh)")
@pytest.mark.parametrize("name", ["an4_shape", "lm_bonus", "peaky_lm", "peaky_minlen", "key_decoder", "primer", "mid_vocab",
                                  "normalize_length", "beam1"])
def test_reference_batch_beam_search_drives_cuda_scorer(name):
    """The strongest drop-in proof: reference BatchBeamSearch + this package's CTCPrefixScorer."""
    from espnet.nets.batch_beam_search import BatchBeamSearch
    from espnet.nets.beam_search import BeamSearch

    g = Golden(name)
    c = g.spec
    V, eos = c["V"], c["eos"]
    bs = BeamSearch(beam_size=c["beam"], weights=g.weights, scorers=_scorers(g["table"], g["table_lm"], V, eos), sos=c["sos"],
                    eos=eos, vocab_size=V, token_list=None, pre_beam_score_key=c["key"],
                    normalize_length=c.get("norm", False), hyp_primer=c.get("primer"))
    bs.__class__ = BatchBeamSearch  # espnet2/bin/asr_inference.py:379
    bs.to(device=DEV, dtype=torch.float32).eval()
    with torch.no_grad():
        nbest = bs(x=torch.from_numpy(g["logp"]).to(DEV), maxlenratio=c["maxr"], minlenratio=c["minr"])
    ref = g.nbest()
    assert len(nbest) == len(ref) >= 1
    for h, (toks, score) in zip(nbest, ref):
        assert h.yseq.tolist() == toks
        assert abs(float(h.score) - score) <= SCORE_ATOL


@pytest.mark.skipif(not _have_ref(), reason="baseline/_ref not installed (scripts/install_ref.sh)")
@pytest.mark.parametrize("graphs", [False, True])
def test_reference_transformer_decoder_under_device_protocol(graphs):
    """The reference's own TransformerDecoder (random-init) as the full scorer of this package's search: its plain
    batch_score protocol (per-hypothesis state lists, one utterance at a time) against DeviceCachedDecoderScorer (all row
    slots of all utterances in one call, memory keys / values per utterance, self-attention key / value caches re-ordered
    on the device, optional CUDA graphs).  Same n-best lists."""
    from espnet2.asr.decoder.transformer_decoder import TransformerDecoder

    from llm_guided_asr_b200 import BatchBeamSearch, CTCPrefixScorer, LengthBonus
    from llm_guided_asr_b200.device_cache import DeviceCachedDecoderScorer

    V, D, beam = 150, 32, 4
    eos = V - 1
    torch.manual_seed(0)
    dec = TransformerDecoder(V, D, attention_heads=4, linear_units=64, num_blocks=3, dropout_rate=0.0, positional_dropout_rate=0.0,
                             self_attention_dropout_rate=0.0, src_attention_dropout_rate=0.0).to(DEV).eval()

    class ProjCTC(torch.nn.Module):
        def __init__(self):
            super().__init__()
            torch.manual_seed(2)
            self.lo = torch.nn.Linear(D, V)

        def log_softmax(self, hs):
            return torch.log_softmax(self.lo(hs), dim=-1)

    ctc = ProjCTC().to(DEV).eval()
    g = torch.Generator().manual_seed(4)
    xs = [torch.randn(T, D, generator=g).to(DEV) for T in (33, 27, 33)]
    weights = dict(decoder=0.7, ctc=0.3, length_bonus=0.2)

    def run(scorer, batch, **kw):
        bs = BatchBeamSearch(scorers=dict(decoder=scorer, ctc=CTCPrefixScorer(ctc=ctc, eos=eos), length_bonus=LengthBonus(V)),
                             weights=weights, beam_size=beam, vocab_size=V, sos=eos, eos=eos, pre_beam_score_key="full", **kw)
        with torch.no_grad():
            return bs.decode_batch(batch, -10, 0.0)

    plain = [run(dec, [x], device_caches=False)[0] for x in xs]  # the reference's own batch_score / state-list protocol
    wrapped = DeviceCachedDecoderScorer(dec, cuda_graphs=graphs)
    for rep in range(2):  # the second batch replays the captured graphs
        both = run(wrapped, xs)
        assert wrapped._fast
        for p, q in zip(plain, both):
            assert len(p) == len(q) >= 1
            for x, y in zip(p, q):
                assert x.yseq.tolist() == y.yseq.tolist()
                assert abs(float(x.score) - float(y.score)) <= SCORE_ATOL
