"""Continuous batching (``lgasr_refill`` / ``BatchBeamSearch.decode_stream``): utterances of different lengths through a
few lock-step slots, every slot re-filled the moment its utterance ends, must give exactly the n-best lists of decoding
each utterance alone (``forward``), whatever the slot count and the order in which the utterances end."""

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _setup(V, beam, with_lm):
    from llm_guided_asr_b200 import BatchBeamSearch, CTCPrefixScorer, LengthBonus
    from llm_guided_asr_b200.testing import synthetic as syn

    eos = V - 1
    _, table, _ = syn.peaky_case(50, 60, V, 9, eos)
    scorers = dict(decoder=syn.PrefixTableScorer(table.to(DEV)), ctc=CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos),
                   length_bonus=LengthBonus(V))
    weights = dict(decoder=0.7, ctc=0.3, length_bonus=0.1)
    if with_lm:
        scorers["lm"] = syn.PrefixTableScorer(syn.random_table(51, table.shape[0], 3, V, scale=1.0).to(DEV))
        weights["lm"] = 0.4
    bs = BatchBeamSearch(scorers=scorers, weights=weights, beam_size=beam, vocab_size=V, sos=eos, eos=eos,
                         pre_beam_score_key="full")
    return bs, eos


def _same(a, b):
    assert len(a) == len(b)
    for x, y in zip(a, b):
        assert x.yseq.tolist() == y.yseq.tolist()
        assert float(x.score) == float(y.score)
        for k in x.scores:
            assert float(x.scores[k]) == float(y.scores[k])


@pytest.mark.parametrize("slots,maxr,with_lm", [(3, 0.0, False), (2, -9.0, True), (5, 0.0, True), (1, 0.0, False)])
def test_decode_stream_equals_single_utterance_decodes(slots, maxr, with_lm):
    from llm_guided_asr_b200.testing import synthetic as syn

    V, beam = 120, 5
    bs, eos = _setup(V, beam, with_lm)
    lens = [60, 31, 44, 72, 25, 58, 39]
    xs = [syn.peaky_case(60 + i, T, V, 4 + (i * 3) % 7, eos)[0].to(DEV) for i, T in enumerate(lens)]
    single, single_steps = [], []
    for x in xs:
        single.append(bs(x, maxr, 0.0))
        single_steps.append(bs.last_steps)
    stream = bs.decode_stream(xs, slots, maxr, 0.0)
    stream_steps = bs.last_steps
    assert len(stream) == len(xs)
    for a, b in zip(single, stream):
        assert len(a) >= 1
        _same(a, b)
    # the slots really overlap: fewer launches than one utterance after the other, at least the per-slot share
    if slots == 1:
        assert stream_steps == sum(single_steps)
    else:
        assert -(-sum(single_steps) // slots) <= stream_steps < sum(single_steps)
    # a handle that was re-filled keeps working for an ordinary batch afterwards
    again = bs.decode_batch(xs[:slots], maxr, 0.0)
    for a, b in zip(single[:slots], again):
        _same(a, b)
