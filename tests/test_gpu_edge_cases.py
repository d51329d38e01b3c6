"""GPU parity on the shapes that exercise the less-travelled paths of k_step, against the oracle
(itself pinned to the reference goldens): long multi-chunk rows with two scorer matrices, a
vocabulary that is not a multiple of 4 (plain-load staging instead of TMA), degenerate rows whose
scores are all equal (pending-list overflow rounds), a constant-only scorer set, the host-buffer
replay against the step-by-step path, and BASELINE-size properties that need no oracle.

Bar: tokens bit-exact, scores within 1e-4 abs (fp32); exact-score ties follow the documented
"higher score, then lower flat index" rule on both sides.
"""

import numpy as np
import pytest
import torch

from helpers import SCORE_ATOL
from oracle.beam_search_oracle import BeamSearchOracle, torch_scorer_fn

pytestmark = pytest.mark.gpu


def _search(scorers, weights, beam, V, eos, key="full"):
    from llm_guided_asr_b200 import BatchBeamSearch

    return BatchBeamSearch(scorers=scorers, weights=weights, beam_size=beam, vocab_size=V, sos=eos, eos=eos,
                           pre_beam_score_key=key)


def _check(nbest, ref, n_min=1):
    assert len(nbest) == len(ref) and len(ref) >= n_min
    for h, r in zip(nbest, ref):
        assert h.yseq.tolist() == r.yseq
        assert abs(float(h.score) - float(r.score)) <= SCORE_ATOL, (float(h.score), float(r.score))


def _case(seed, T, V, beam, ntok, weights, with_lm, key="full", maxr=0.0):
    from llm_guided_asr_b200 import CTCPrefixScorer, LengthBonus
    from llm_guided_asr_b200.testing import synthetic as syn

    dev = torch.device("cuda:0")
    eos = V - 1
    logp, table, _ = syn.peaky_case(seed, T, V, ntok, eos)
    table_lm = syn.random_table(seed + 1, table.shape[0], 3, V, scale=1.0)
    scorers = dict(decoder=syn.PrefixTableScorer(table.to(dev)), ctc=CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos),
                   length_bonus=LengthBonus(V))
    full = dict(decoder=torch_scorer_fn(syn.PrefixTableScorer(table)),
                length_bonus=lambda ys: np.ones((len(ys), V), dtype=np.float32))
    if with_lm:
        scorers["lm"] = syn.PrefixTableScorer(table_lm.to(dev))
        full["lm"] = torch_scorer_fn(syn.PrefixTableScorer(table_lm))
    full = {k: v for k, v in full.items() if weights.get(k, 0) != 0}
    bs = _search(scorers, weights, beam, V, eos, key)
    nbest = bs(logp.to(dev), maxr, 0.0)
    orc = BeamSearchOracle(full, weights, beam, V, eos, eos, pre_beam_score_key=key)
    ref = orc.forward(logp.numpy(), maxr, 0.0)
    return nbest, ref


@pytest.mark.parametrize("no_cluster", ["0", "1"])
def test_large_vocab_two_matrices_multi_chunk(monkeypatch, no_cluster):
    monkeypatch.setenv("LGASR_NO_CLUSTER", no_cluster)
    _large_vocab()


def _large_vocab():
    """OWSM-like: V = 20000, beam 5, pre-beam 7, decoder + LM matrices -> 2 stages, 14 chunks per row."""
    nbest, ref = _case(21, 40, 20000, 5, 6, dict(decoder=0.7, ctc=0.3, lm=0.5, length_bonus=0.2), True)
    _check(nbest, ref)


def test_vocab_not_multiple_of_four_uses_plain_staging():
    nbest, ref = _case(22, 48, 1001, 4, 7, dict(decoder=0.6, ctc=0.4, lm=0.3), True)
    _check(nbest, ref)


@pytest.mark.parametrize("no_cluster", ["0", "1"])
def test_single_matrix_long_row_one_chunk(monkeypatch, no_cluster):
    """The bench shape in small: one scorer matrix, V = 5000 (one 20 KB chunk), beam 10; with the records
    exchanged through distributed shared memory (cluster launch) and through L2 (ticket)."""
    monkeypatch.setenv("LGASR_NO_CLUSTER", no_cluster)
    nbest, ref = _case(23, 60, 5000, 10, 8, dict(decoder=0.7, ctc=0.3), False, maxr=-12)
    _check(nbest, ref)


def test_long_peaky_utterance_full_length():
    """configs[1] length (T = 374 frames) with peaky posteriors: the r values reach magnitudes where one fp32
    ulp is ~3e-5, the regime the MUFU log-add-exp of the scans has to survive; ends through end_detect."""
    nbest, ref = _case(25, 374, 800, 10, 28, dict(decoder=0.7, ctc=0.3, lm=0.4, length_bonus=0.5), True)
    _check(nbest, ref, n_min=5)


def test_owsm_shape_vocab_50000_two_matrices():
    """configs[3] shape: T = 749 frames, V = 50 000, beam 5 (pre-beam 7), decoder + LM matrices: rows of
    200 KB stream through the two-stage ring in many chunks; 3 KB state/posterior rows."""
    nbest, ref = _case(26, 749, 50000, 5, 8, dict(decoder=0.7, ctc=0.3, lm=0.5), True, maxr=-10)
    _check(nbest, ref)


def test_pre_beam_key_on_one_scorer():
    nbest, ref = _case(24, 50, 3000, 6, 6, dict(decoder=0.7, ctc=0.3, lm=0.4), True, key="decoder")
    _check(nbest, ref)


def test_all_equal_rows_overflow_rounds():
    """Every decoder score equal: the pre-beam set is decided purely by the index tie-break and the
    pending list (256 entries) overflows on every row -> the multi-round path."""
    from llm_guided_asr_b200 import CTCPrefixScorer
    from llm_guided_asr_b200.testing import synthetic as syn

    dev = torch.device("cuda:0")
    T, V, beam = 30, 1200, 4
    eos = V - 1
    logp = syn.random_logp(31, T, V)

    class Flat(syn.BatchScorerInterface):
        def score(self, y, state, x):
            return torch.full((V,), -7.5, device=y.device), None

        def batch_score(self, ys, states, xs):
            return torch.full((ys.shape[0], V), -7.5, device=ys.device), [None] * ys.shape[0]

    weights = dict(decoder=0.7, ctc=0.3)
    bs = _search(dict(decoder=Flat(), ctc=CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos)), weights, beam, V, eos)
    nbest = bs(logp.to(dev), -6, 0.0)
    full = dict(decoder=lambda ys: np.full((len(ys), V), -7.5, dtype=np.float32))
    ref = BeamSearchOracle(full, weights, beam, V, eos, eos, pre_beam_score_key="full").forward(logp.numpy(), -6, 0.0)
    _check(nbest, ref)


@pytest.mark.parametrize("no_cluster", ["0", "1"])
def test_replay_host_equals_stepping_at_full_vocab_width(monkeypatch, no_cluster):
    monkeypatch.setenv("LGASR_NO_CLUSTER", no_cluster)
    _replay_host_equals_stepping()


def _replay_host_equals_stepping():
    """Whole-batch replay from host buffers (graph + PDL launches, double-buffered copies) must give
    exactly what position-by-position stepping gives: B = 8 utterances, ragged lengths, V = 5000."""
    from llm_guided_asr_b200.engine import FusedStepEngine
    from llm_guided_asr_b200.testing import synthetic as syn

    dev = torch.device("cuda:0")
    B, T, V, W, S, L = 8, 120, 5000, 10, 15, 12
    eos = V - 1
    lens = [120, 97, 120, 64, 33, 120, 101, 88]
    logp = torch.stack([syn.random_logp(40 + b, T, V) for b in range(B)])
    g = torch.Generator().manual_seed(3)
    tables = torch.log_softmax(2.0 * torch.randn(L, B * W, V, generator=g), dim=-1)
    eng = FusedStepEngine(B, T, V, W, S, ["decoder"], "full", True, 0.3, eos, eos, L, 1, False, device=dev)
    mx, mn = [L] * B, [0] * B

    def collect():
        return [tuple(map(lambda a: np.asarray(a).tolist(), eng.gather_ended(b)[:2])) for b in range(B)]

    d_logp, d_tab = logp.to(dev), tables.to(dev)
    eng.set_posteriors(d_logp, lens)
    eng.begin(mx, mn, [eos])
    for i in range(L):
        eng.step([d_tab[i]], [0.0], [0.7])
        eng.poll(with_parents=False)
    stepped = collect()
    # lgasr_gather_best == the first maximum of gather_ended, for every utterance
    best = eng.gather_best()
    for b, (toks, score) in enumerate(stepped):
        j = int(np.argmax(np.asarray(score, dtype=np.float32)))
        assert best[b][0] == toks[j] and best[b][1] == np.float32(score[j])
    for _ in range(2):  # second pass replays the cached graph
        eng.set_posteriors(d_logp, lens)
        eng.begin(mx, mn, [eos])
        eng.replay([[d_tab[i]] for i in range(L)], [0.0], [0.7], ld=V)
        assert collect() == stepped
    h_logp, h_tab = logp.pin_memory(), tables.pin_memory()
    eng.replay_host(h_logp, lens, mx, mn, [eos], [[h_tab[i]] for i in range(L)], [0.0], [0.7], ld=V)
    assert collect() == stepped
    # size-independent properties: every ended hypothesis starts with sos, ends with eos, has L + 2 tokens
    # (maxlen reached), and the scores of one utterance come out sorted by the final top-k
    for toks, score in stepped:
        assert len(toks) == W
        for t in toks:
            assert t[0] == eos and t[-1] == eos and len(t) == L + 2
        assert all(np.isfinite(score))
    eng.close()


@pytest.mark.parametrize(
    "V,names,key,use_ctc,consts,flags",
    [
        (20000, ["decoder", "lm"], "full", True, None, "0"),  # two matrices, 2-stage ring, many chunks
        (20000, ["decoder", "lm"], "full", True, None, "1"),  # ... with the per-utterance flag hand-off
        (1001, ["decoder", "lm"], "full", True, None, "0"),  # vocabulary not a multiple of 4: plain-load staging
        (3000, ["decoder", "lm"], "decoder", True, None, "0"),  # pre-beam on one scorer
        (3000, ["decoder", "length_bonus"], "full", True, 1.0, "0"),  # a constant term
        (2000, ["decoder"], None, False, None, "0"),  # no CTC scorer: top-beam per hypothesis
    ],
)
def test_replay_instances_equal_stepping_on_every_staging_path(monkeypatch, V, names, key, use_ctc, consts, flags):
    """The replay (CUDA graph + programmatic dependent launch: the SEQ instance of k_step, selection before
    griddepcontrol.wait, four CTC warps) must reproduce position-by-position stepping (the side-by-side
    instance) bit for bit on every staging path; ragged lengths, an utterance that finishes early."""
    from llm_guided_asr_b200.engine import FusedStepEngine
    from llm_guided_asr_b200.testing import synthetic as syn

    monkeypatch.setenv("LGASR_FLAGS", flags)
    dev = torch.device("cuda:0")
    B, T, W, L = 5, 70, 6, 10
    eos = V - 1
    lens = [70, 41, 70, 9, 55]
    logp = torch.stack([syn.random_logp(60 + b, T, V) for b in range(B)]).to(dev)
    g = torch.Generator().manual_seed(V)
    nmat = sum(1 for n in names if n != "length_bonus")
    tabs = torch.log_softmax(2.0 * torch.randn(L, nmat, B * W, V, generator=g), dim=-1).to(dev)
    wts = [0.7, 0.4][: len(names)]
    cst = [0.0] * len(names)
    if consts is not None:
        cst[names.index("length_bonus")] = consts
    eng = FusedStepEngine(B, T, V, W, int(1.5 * W), names, key, use_ctc, 0.3, eos, eos, L, 1, True, device=dev)
    mx, mn = [L, L, L, 3, L], [0] * B  # utterance 3 stops after 3 positions

    def mats(i):
        out, m = [], 0
        for n in names:
            if n == "length_bonus":
                out.append(None)
            else:
                out.append(tabs[i, m])
                m += 1
        return out

    def collect():
        return [tuple(np.asarray(a).tolist() for a in eng.gather_ended(b)[:2]) for b in range(B)]

    eng.set_posteriors(logp if use_ctc else None, lens)
    eng.begin(mx, mn, [eos])
    for i in range(L):
        eng.step(mats(i), cst, wts)
        if eng.poll(with_parents=False)[4] == 0:
            break
    stepped = collect()
    assert any(len(t) for t, _ in stepped)
    for _ in range(2):
        eng.set_posteriors(logp if use_ctc else None, lens)
        eng.begin(mx, mn, [eos])
        eng.replay([mats(i) for i in range(L)], cst, wts, ld=V)
        assert collect() == stepped
    eng.close()


@pytest.mark.parametrize("beam,V,T", [(20, 600, 64), (40, 2500, 120), (16, 900, 50), (17, 900, 50), (85, 3000, 80)])
def test_wide_beams(beam, V, T):
    """Beams above the bench shape: beam 16 is the largest cluster, 17 the first ticket-only width; from beam 20 on
    the selection sees more than 512 records (rounds of block arg-best instead of rank counting) and the
    candidates' posterior rows no longer fit one gather group."""
    nbest, ref = _case(70 + beam, T, V, beam, 9, dict(decoder=0.7, ctc=0.3, lm=0.3, length_bonus=0.1), True, maxr=-14)
    _check(nbest, ref, n_min=1)


def test_pre_beam_wider_than_the_limit_is_refused_loudly():
    """int(1.5 * 86) + 1 = 130 records per hypothesis exceed LGASR_MAX_PREBEAM: LGASR_EINVAL, no silent fallback."""
    from llm_guided_asr_b200._lib import LgasrError

    with pytest.raises(LgasrError):
        _case(99, 40, 3000, 86, 5, dict(decoder=0.7, ctc=0.3), False, maxr=-4)
