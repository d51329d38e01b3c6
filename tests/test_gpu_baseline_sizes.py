"""GPU parity at BASELINE.json sizes and on the launch variants the small goldens never reach.

* configs[4] / configs[2] shape: T = 374, V = 5000, beam 10, TWO full-scorer matrices, B = 64 (one wave, cluster
  exchange), 128 and 512 (several waves, ticket exchange): the stepping launches and the graph replay must agree on
  every utterance, and a sample of utterances is re-decoded by the oracle (itself pinned to the reference goldens).
* OWSM shape (T = 749, V = 50 000, beam 5, two matrices) at B = 32 with ragged utterance lengths.

Inputs are generated on the device (the full-size tensors take GBs); the sampled utterances are copied back for the
oracle.  Bar: tokens bit-exact, scores within 1e-4 abs (fp32).
"""

import numpy as np
import pytest
import torch

from helpers import SCORE_ATOL
from oracle.beam_search_oracle import BeamSearchOracle

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _make(B, T, V, W, K, L, seed, xlens=None):
    g = torch.Generator(device=DEV).manual_seed(seed)
    logp = torch.log_softmax(3.0 * torch.randn(B, T, V, device=DEV, generator=g), dim=-1)
    tabs = [torch.log_softmax(2.0 * torch.randn(L, B * W, V, device=DEV, generator=g), dim=-1) for _ in range(K)]
    return logp, tabs


def _engine(B, T, V, W, S, K, L, end_detect=False):
    from llm_guided_asr_b200.engine import FusedStepEngine

    names = ["decoder", "lm"][:K]
    return FusedStepEngine(B, T, V, W, S, names, "full", True, 0.3, V - 1, V - 1, L, 1, end_detect, device=DEV)


def _oracle_utt(logp_b, tabs_b, W, V, L, wts, T_b):
    """tabs_b[k]: (L, W, V) numpy; the oracle sees the same rows the engine's row slots see."""
    names = ["decoder", "lm"][: len(tabs_b)]
    full = {n: (lambda ys, t=t: t[ys.shape[1] - 1, : len(ys)]) for n, t in zip(names, tabs_b)}
    weights = dict(zip(names, wts))
    weights["ctc"] = 0.3
    orc = BeamSearchOracle(full, weights, W, V, V - 1, V - 1, pre_beam_score_key="full")
    return orc.forward(logp_b[:T_b], -L, 0.0)


def _run_shape(B, T, V, W, S, K, L, sample, xlens=None, seed=5):
    logp, tabs = _make(B, T, V, W, K, L, seed)
    wts = [0.7, 0.5][:K]
    consts = [0.0] * K
    xl = [T] * B if xlens is None else list(xlens)
    eng = _engine(B, T, V, W, S, K, L)
    info = eng.launch_info()
    mats = [[t[i] for t in tabs] for i in range(L)]
    # stepping launches (what a decoder in the loop uses)
    eng.set_posteriors(logp, xl)
    eng.begin([L] * B, [0] * B, [V - 1])
    for i in range(L):
        eng.step(mats[i], consts, wts)
    step_best = eng.gather_best()
    step_nbest = {b: eng.gather_ended(b) for b in sample}
    # graph + programmatic-dependent-launch replay
    eng.begin([L] * B, [0] * B, [V - 1])
    eng.replay(mats, consts, wts, ld=V)
    rep_best = eng.gather_best()
    assert rep_best == step_best, "replay and stepping launches disagree"
    assert all(len(t) == L + 2 for t, _ in step_best)
    # oracle on the sampled utterances: whole n-best lists
    for b in sample:
        ref = _oracle_utt(logp[b].cpu().numpy(), [t[:, b * W : (b + 1) * W].cpu().numpy() for t in tabs], W, V, L, wts, xl[b])
        toks, score, _ = step_nbest[b]
        order = sorted(range(len(toks)), key=lambda e: -float(score[e]))
        assert len(order) == len(ref) >= 1
        for e, r in zip(order, ref):
            assert toks[e] == r.yseq, (b, toks[e], r.yseq)
            assert abs(float(score[e]) - float(r.score)) <= SCORE_ATOL
    eng.close()
    return info


@pytest.mark.parametrize("B", [64, 128, 512])
def test_conformer_lm_shape_two_matrices(B):
    """configs[2] (B = 64, one wave) and configs[4] per-GPU batches (B = 128 / 512: several waves, ticket exchange)."""
    rng = np.random.default_rng(B)
    sample = sorted(int(b) for b in rng.choice(B, size=8, replace=False))
    info = _run_shape(B, 374, 5000, 10, 15, 2, 12, sample)
    if B >= 128:
        assert not info["cluster"], "several waves must take the ticket exchange"


def test_conformer_shape_one_matrix_512():
    """configs[1] rows at configs[4]'s per-GPU batch: one matrix, several waves."""
    _run_shape(512, 374, 5000, 10, 15, 1, 12, [0, 77, 300, 511])


def test_owsm_shape_ragged_batch_32():
    """configs[3]: T = 749, V = 50 000, beam 5, pre-beam 7, two matrices, 32 utterances of different lengths."""
    B, T = 32, 749
    rng = np.random.default_rng(3)
    xlens = [int(v) for v in rng.integers(400, T + 1, size=B)]
    xlens[0] = T
    _run_shape(B, T, 50000, 5, 7, 2, 6, [0, 5, 17, 31], xlens=xlens)
