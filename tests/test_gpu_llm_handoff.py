"""The LLM / LM logits hand-off on the GPU.

* ``lgasr_step_ex``: a bf16 logits matrix plus one logsumexp per row (``lgasr_row_logsumexp``) must decode exactly like
  the fp32 log-probability matrix ``lgasr_log_softmax_rows`` makes of the same logits, and like the oracle fed by
  ``torch.log_softmax(logits.float())``; short rows, long rows (several load groups behind the L2 prefetch), mixed with a
  plain fp32 decoder matrix and a constant term.
* ``DevicePrefixLMScorer`` around a random-init GPT-2 (``hf_lm.gpt2_small`` scaled down): device protocol (all utterances
  in one call, fused log-softmax / fused normaliser / CUDA graphs) against the plain scorer protocol and the oracle.
* ``DeviceLLMGuidedDecoderScorer`` around the reference's own ``LLMGuidedTransformerDecoder`` (``baseline/_ref``) with a
  stand-in LLM: device protocol against the reference object's ``batch_score`` under the same search.

Bar: tokens bit-exact; scores within 1e-4 abs (fp32), exact where both sides run the same arithmetic.
"""

import numpy as np
import pytest
import torch

from helpers import SCORE_ATOL
from oracle.beam_search_oracle import BeamSearchOracle, torch_scorer_fn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _nbest_engine(eng, B):
    out = []
    for b in range(B):
        toks, score, _ = eng.gather_ended(b)
        order = sorted(range(len(toks)), key=lambda e: -float(score[e]))
        out.append([(toks[e], float(score[e])) for e in order])
    return out


@pytest.mark.parametrize("same_kernel", ["1", "0"])
@pytest.mark.parametrize("V,T,B,W,S", [(1000, 48, 3, 4, 6), (32000, 40, 2, 5, 7)])
def test_step_ex_bf16_logits_and_normaliser(V, T, B, W, S, same_kernel, monkeypatch):
    # typed term matrices run k_step2; plain fp32 ones run k_step unless LGASR_V2=1 sends them through k_step2 too:
    # with the same kernel on both routes the n-best lists must be bit-identical, otherwise within the tolerance
    monkeypatch.setenv("LGASR_V2", same_kernel)
    from llm_guided_asr_b200.device_cache import log_softmax_rows, row_logsumexp
    from llm_guided_asr_b200.engine import FusedStepEngine
    from llm_guided_asr_b200.testing import synthetic as syn

    L = 8
    eos = V - 1
    g = torch.Generator().manual_seed(V + T)
    logp = torch.stack([syn.random_logp(100 + b, T, V) for b in range(B)])
    dec = torch.log_softmax(2.0 * torch.randn(L, B * W, V, generator=g), dim=-1).to(DEV)
    llm_logits = (3.0 * torch.randn(L, B * W, V, generator=g)).to(torch.bfloat16).to(DEV)  # what an LLM head emits
    wts = [0.7, 0.4, 0.1]
    consts = [0.0, 0.0, 1.0]  # decoder, llm, length bonus

    def decode(typed):
        eng = FusedStepEngine(B, T, V, W, S, ["decoder", "llm", "length_bonus"], "full", True, 0.3, eos, eos, L, 1, False, device=DEV)
        eng.set_posteriors(logp.to(DEV), [T] * B)
        eng.begin([L] * B, [0] * B, [eos])
        for i in range(L):
            if typed:  # logits (bf16) + normaliser: the row is read once, at 2 bytes per entry
                lse = row_logsumexp(llm_logits[i])
                eng.step([dec[i], llm_logits[i], None], consts, wts, [None, lse, None])
            else:  # the two-pass route: fp32 log-probabilities written first
                eng.step([dec[i], log_softmax_rows(llm_logits[i]), None], consts, wts)
        res = _nbest_engine(eng, B)
        eng.close()
        return res

    a, c = decode(True), decode(False)
    if same_kernel == "1":
        assert a == c, "logits + normaliser and the fp32 log-probability matrix must give identical n-best lists"
    for ua, uc in zip(a, c):
        assert len(ua) == len(uc)
        for (ta, sa), (tc, sc_) in zip(ua, uc):
            assert ta == tc
            assert abs(sa - sc_) <= SCORE_ATOL
    # the oracle sees what the reference would: log_softmax of the fp32-converted logits (transformer_decoder.py:212)
    llm_lp = torch.log_softmax(llm_logits.float(), dim=-1).cpu().numpy()
    dec_np = dec.cpu().numpy()
    for b in range(B):
        full = dict(decoder=lambda ys, b=b: dec_np[ys.shape[1] - 1, b * W : b * W + len(ys)],
                    llm=lambda ys, b=b: llm_lp[ys.shape[1] - 1, b * W : b * W + len(ys)],
                    length_bonus=lambda ys: np.ones((len(ys), V), dtype=np.float32))
        orc = BeamSearchOracle(full, dict(decoder=0.7, llm=0.4, length_bonus=0.1, ctc=0.3), W, V, eos, eos, pre_beam_score_key="full")
        ref = orc.forward(logp[b].numpy(), -L, 0.0)
        assert len(ref) == len(a[b]) >= 1
        for (toks, sc), r in zip(a[b], ref):
            assert toks == r.yseq
            assert abs(sc - float(r.score)) <= SCORE_ATOL


def _small_case(seed, T, V, ntok):
    from llm_guided_asr_b200.testing import synthetic as syn

    eos = V - 1
    logp, table, _ = syn.peaky_case(seed, T, V, ntok, eos)
    return logp, table, eos


def _search(scorers, weights, beam, V, eos):
    from llm_guided_asr_b200 import BatchBeamSearch

    return BatchBeamSearch(scorers=scorers, weights=weights, beam_size=beam, vocab_size=V, sos=eos, eos=eos,
                           pre_beam_score_key="full")


@pytest.mark.parametrize("mode", ["softmax", "normaliser", "graphs"])
def test_device_prefix_lm_scorer_gpt2(mode):
    from llm_guided_asr_b200 import CTCPrefixScorer
    from llm_guided_asr_b200.device_cache import DevicePrefixLMScorer
    from llm_guided_asr_b200.hf_lm import gpt2_small
    from llm_guided_asr_b200.testing import synthetic as syn

    V, beam, T = 200, 5, 40
    cases = [_small_case(31 + i, T - 4 * i, V, 6) for i in range(2)]
    eos = V - 1
    lm = gpt2_small(V, n_positions=64, n_layer=2, n_embd=64, n_head=4, seed=3).to(DEV)
    weights = dict(decoder=0.7, ctc=0.3, lm=0.4)

    def scorers(lm_scorer, table):
        return dict(decoder=syn.PrefixTableScorer(table.to(DEV)), ctc=CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos), lm=lm_scorer)

    wrapped = DevicePrefixLMScorer(lm, fused_normaliser=(mode == "normaliser"), cuda_graphs=(mode == "graphs"))
    # the tables differ per utterance: decode one utterance at a time through both protocols, then the two together
    with torch.no_grad():
        plain = [_search(scorers(lm, tb), weights, beam, V, eos)(lp.to(DEV), -10, 0.0) for lp, tb, _ in cases]
        dev1 = [_search(scorers(wrapped, tb), weights, beam, V, eos)(lp.to(DEV), -10, 0.0) for lp, tb, _ in cases]
    for p, q in zip(plain, dev1):
        assert len(p) == len(q) >= 1
        for x, y in zip(p, q):
            assert x.yseq.tolist() == y.yseq.tolist()
            assert abs(float(x.score) - float(y.score)) <= SCORE_ATOL
    # oracle (the reference's arithmetic, fed by the same network on the CPU)
    lm_cpu = gpt2_small(V, n_positions=64, n_layer=2, n_embd=64, n_head=4, seed=3)
    for (lp, tb, _), q in zip(cases, dev1):
        full = dict(decoder=torch_scorer_fn(syn.PrefixTableScorer(tb)), lm=torch_scorer_fn(lm_cpu))
        ref = BeamSearchOracle(full, weights, beam, V, eos, eos, pre_beam_score_key="full").forward(lp.numpy(), -10, 0.0)
        assert len(ref) == len(q)
        for r, y in zip(ref, q):
            assert r.yseq == y.yseq.tolist()
            assert abs(float(r.score) - float(y.score)) <= 5e-4  # a 2-layer network on two devices: 5e-4


def _have_ref():
    try:
        import espnet2.asr.decoder.transformer_decoder  # noqa: F401

        return True
    except Exception:
        return False


@pytest.mark.skipif(not _have_ref(), reason="baseline/_ref not installed (scripts/install_ref.sh)")
@pytest.mark.parametrize("fused_normaliser", [False, True])
def test_device_llm_guided_decoder_scorer_against_reference_class(fused_normaliser):
    """The reference's own LLMGuidedTransformerDecoder (random-init attention layers, a stand-in LLM) under this
    package's search: its plain batch_score (one utterance at a time) against the device protocol wrapper."""
    from espnet2.asr.decoder.transformer_decoder import LLMGuidedTransformerDecoder

    from llm_guided_asr_b200 import CTCPrefixScorer
    from llm_guided_asr_b200.device_cache import DeviceLLMGuidedDecoderScorer

    V, D, DL, beam = 120, 32, 24, 4
    eos = V - 1

    class StubLLM(torch.nn.Module):
        """forward_inference(hyps, hyps_lengths, tgt, tgt_lengths) -> (hidden states of the response tokens, lengths):
        a causal mix of token embeddings conditioned on the prompt (the CTC hypothesis)."""

        start_of_response_token_id = 1

        def __init__(self):
            super().__init__()
            torch.manual_seed(5)
            self.emb = torch.nn.Embedding(V, DL)
            self.mix = torch.nn.Linear(DL, DL)

        def forward_inference(self, hyps, hyps_lengths, tgt, tgt_lengths):
            prompt = torch.stack([self.emb(h.clamp(0, V - 1)).mean(0) if h.numel() else self.emb.weight[0] * 0 for h in hyps])
            x = self.emb(tgt)
            csum = torch.cumsum(x, dim=1) / torch.arange(1, x.shape[1] + 1, device=x.device)[None, :, None]
            return torch.tanh(self.mix(csum + prompt[:, None, :])), tgt_lengths

    class StubCTC(torch.nn.Module):
        def __init__(self):
            super().__init__()
            torch.manual_seed(6)
            self.lo = torch.nn.Linear(D, V)

        def log_softmax(self, hs):
            return torch.log_softmax(self.lo(hs), dim=-1)

        def argmax(self, hs):
            return torch.argmax(self.lo(hs), dim=2)

    torch.manual_seed(0)
    dec = LLMGuidedTransformerDecoder(vocab_size=V, encoder_output_size=D, attention_heads=4, linear_units=64, num_blocks=2,
                                      dropout_rate=0.0, positional_dropout_rate=0.0, self_attention_dropout_rate=0.0,
                                      src_attention_dropout_rate=0.0)
    dec.llm = StubLLM()
    dec.embed = torch.nn.Linear(DL, D)
    dec.ctc = StubCTC()
    dec = dec.to(DEV).eval()
    ctc = dec.ctc
    g = torch.Generator().manual_seed(9)
    xs = [torch.randn(T, D, generator=g).to(DEV) for T in (30, 24)]
    weights = dict(decoder=0.7, ctc=0.3)

    def run(scorer, batch):
        from llm_guided_asr_b200 import BatchBeamSearch

        bs = BatchBeamSearch(scorers=dict(decoder=scorer, ctc=CTCPrefixScorer(ctc=ctc, eos=eos)), weights=weights,
                             beam_size=beam, vocab_size=V, sos=eos, eos=eos, pre_beam_score_key="full")
        with torch.no_grad():
            return bs.decode_batch(batch, -8, 0.0)

    plain = [run(dec, [x])[0] for x in xs]  # the reference protocol handles one utterance at a time (self.hyp)
    wrapped = DeviceLLMGuidedDecoderScorer(dec, fused_normaliser=fused_normaliser)
    both = run(wrapped, xs)
    for p, q in zip(plain, both):
        assert len(p) == len(q) >= 1
        for x, y in zip(p, q):
            assert x.yseq.tolist() == y.yseq.tolist()
            assert abs(float(x.score) - float(y.score)) <= SCORE_ATOL
