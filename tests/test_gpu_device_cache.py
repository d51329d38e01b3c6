"""GPU: the kernels either side of the fused step, through the C ABI.

* ``lgasr_log_softmax_rows`` against ``torch.log_softmax`` in fp32 (tolerance 2e-5 abs: same formula, another
  summation order), fp32 and bf16 logits, vocabularies from 30 to 128 256 (cluster of 8 CTAs per row), strided rows;
* ``lgasr_reorder_rows`` bit-exact against ``index_select`` with the parents ``lgasr_poll`` reports, rows of idle
  slots untouched, 16-byte and 4-byte paths, differing source / destination strides;
* a decode with the decoder caches kept on the device (``DeviceCachedDecoderScorer``) against (i) the same search
  with per-hypothesis host-side states (the reference's protocol) and (ii) the oracle search fed by a cache-free
  full forward of the same decoder on the CPU: tokens equal, scores within 1e-4.
"""

import numpy as np
import pytest
import torch

from helpers import SCORE_ATOL
from oracle.beam_search_oracle import BeamSearchOracle

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.mark.parametrize("V", [30, 1001, 5000, 16392, 50000, 128256])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_log_softmax_rows_matches_torch(V, dtype):
    from llm_guided_asr_b200.device_cache import log_softmax_rows

    g = torch.Generator().manual_seed(V)
    rows = 13
    x = (4.0 * torch.randn(rows, V, generator=g)).to(dtype).to(DEV)
    x[3, V // 2] = 40.0  # a dominant entry
    x[5] = -3.25  # a constant row
    want = torch.log_softmax(x.float(), dim=-1)
    got = log_softmax_rows(x)
    assert got.dtype == torch.float32 and got.shape == (rows, V)
    assert float((got - want).abs().max()) <= 2e-5
    assert float((torch.logsumexp(got, dim=-1)).abs().max()) <= 2e-5


def test_log_softmax_rows_strided_input_and_output():
    from llm_guided_asr_b200.device_cache import log_softmax_rows

    g = torch.Generator().manual_seed(5)
    big = torch.randn(9, 1300, generator=g).to(DEV)
    x = big[:, 3:1203]  # unaligned start, ld 1300
    out_big = torch.full((9, 1400), 7.0, device=DEV)
    out = out_big[:, :1200]
    log_softmax_rows(x, out)
    assert float((out - torch.log_softmax(x, dim=-1)).abs().max()) <= 2e-5
    assert bool((out_big[:, 1200:] == 7.0).all())


def _stepped_engine(B, W, V, seed=0):
    from llm_guided_asr_b200.engine import FusedStepEngine
    from llm_guided_asr_b200.testing import synthetic as syn

    T, L = 40, 6
    eos = V - 1
    eng = FusedStepEngine(B, T, V, W, int(1.5 * W), ["decoder"], "full", True, 0.3, eos, eos, L, 1, False, device=torch.device(DEV))
    logp = torch.stack([syn.random_logp(seed + b, T, V) for b in range(B)]).to(DEV)
    eng.set_posteriors(logp, [T] * B)
    eng.begin([L] * B, [0] * B, [eos])
    g = torch.Generator().manual_seed(seed)
    for i in range(3):
        eng.step([torch.log_softmax(2 * torch.randn(B * W, V, generator=g), -1).to(DEV)], [0.0], [0.7])
    return eng


@pytest.mark.parametrize("elem,width", [(torch.float32, 64), (torch.float32, 3), (torch.bfloat16, 40), (torch.float32, 4096)])
def test_reorder_rows_equals_index_select(elem, width):
    B, W, V = 5, 4, 300
    eng = _stepped_engine(B, W, V)
    nrun, done, par, _tok, _ = eng.poll(with_parents=True)
    nl, cap, L = 3, 9, 5
    g = torch.Generator().manual_seed(1)
    src = torch.randn(nl, B * W, L, width, generator=g).to(elem).to(DEV)
    dst = torch.full((nl, B * W, cap, width), -1.0, dtype=elem, device=DEV)
    eng.reorder_rows(dst, src, row_elems=L * width)
    want = torch.full_like(dst, -1.0)
    for b in range(B):
        for j in range(int(nrun[b])):
            want[:, b * W + j, :L] = src[:, b * W + int(par[b, j])]
    assert torch.equal(dst, want)
    assert int(nrun.sum()) > 0
    # single-layer form, partial rows
    dst2 = torch.zeros(B * W, cap, width, dtype=elem, device=DEV)
    if (2 * width * src.element_size()) % 4 == 0:
        eng.reorder_rows(dst2, src[1], row_elems=2 * width)
        want2 = torch.zeros_like(dst2)
        for b in range(B):
            for j in range(int(nrun[b])):
                want2[b * W + j, :2] = src[1, b * W + int(par[b, j]), :2]
        assert torch.equal(dst2, want2)
    with pytest.raises(RuntimeError):
        eng.reorder_rows(src, src)
    eng.close()


def _full_forward_fn(dec, memory):
    """Oracle full scorer: cache-free forward of the decoder over the whole prefix, CPU."""

    def fn(ys):
        ys = torch.as_tensor(np.asarray(ys))
        n, L = ys.shape
        mask = torch.ones(L, L, dtype=torch.bool).tril().unsqueeze(0)
        with torch.no_grad():
            logp, _ = dec.forward_one_step(ys, mask, memory.unsqueeze(0).expand(n, *memory.shape), cache=None)
        return logp.numpy()

    return fn


@pytest.mark.parametrize("fused_lsm", [True, False])
def test_device_cached_decoder_decode(fused_lsm):
    from llm_guided_asr_b200 import BatchBeamSearch, CTCPrefixScorer, LengthBonus
    from llm_guided_asr_b200.testing import synthetic as syn
    from llm_guided_asr_b200.device_cache import DeviceCachedDecoderScorer

    V, D, beam = 120, 32, 4
    eos = V - 1
    lens = [37, 52, 29]
    g = torch.Generator().manual_seed(11)
    encs = [torch.randn(t, D, generator=g) for t in lens]

    class ProjCTC(torch.nn.Module):  # CTC head: linear + log_softmax (espnet2/asr/ctc.py:196-204)
        def __init__(self):
            super().__init__()
            with torch.random.fork_rng(devices=[]):
                torch.manual_seed(3)
                self.lo = torch.nn.Linear(D, V)

        def log_softmax(self, hs):
            return torch.log_softmax(3.0 * self.lo(hs), dim=-1)

    dec_cpu = syn.TinyDecoder(V, D=D, seed=2).eval()
    ctc_cpu = ProjCTC().eval()
    dec, ctc = syn.TinyDecoder(V, D=D, seed=2).eval().to(DEV), ProjCTC().eval().to(DEV)
    weights = dict(decoder=0.7, ctc=0.3, length_bonus=0.3)

    def search(decoder_scorer, **kw):
        return BatchBeamSearch(scorers=dict(decoder=decoder_scorer, ctc=CTCPrefixScorer(ctc=ctc, eos=eos), length_bonus=LengthBonus(V)),
                               weights=weights, beam_size=beam, vocab_size=V, sos=eos, eos=eos, pre_beam_score_key="full", **kw)

    xs = [e.to(DEV) for e in encs]
    got = search(DeviceCachedDecoderScorer(dec, fused_log_softmax=fused_lsm)).decode_batch(xs, -14, 0.0)
    host = search(dec, device_caches=False).decode_batch(xs, -14, 0.0)  # per-hypothesis host states (the reference's protocol)
    auto = search(dec)  # the bare decoder handed over as it is: wrapped by the search itself
    assert isinstance(auto.full_scorers["decoder"], DeviceCachedDecoderScorer) and auto.nn_dict["decoder"].net is dec
    for u, v in zip(auto.decode_batch(xs, -14, 0.0), got):
        assert [h.yseq.tolist() for h in u] == [h.yseq.tolist() for h in v]
        assert all(abs(float(a.score) - float(b.score)) <= SCORE_ATOL for a, b in zip(u, v))
    for b in range(len(lens)):
        assert len(got[b]) == len(host[b]) >= 1
        for h, r in zip(got[b], host[b]):
            assert h.yseq.tolist() == r.yseq.tolist()
            assert abs(float(h.score) - float(r.score)) <= SCORE_ATOL
        with torch.no_grad():
            logp = ctc_cpu.log_softmax(encs[b].unsqueeze(0))[0].numpy()
        full = dict(decoder=_full_forward_fn(dec_cpu, encs[b]), length_bonus=lambda ys: np.ones((len(ys), V), dtype=np.float32))
        ref = BeamSearchOracle(full, weights, beam, V, eos, eos, pre_beam_score_key="full").forward(logp, -14, 0.0)
        assert len(ref) == len(got[b])
        for h, r in zip(got[b], ref):
            assert h.yseq.tolist() == r.yseq
            assert abs(float(h.score) - float(r.score)) <= SCORE_ATOL


def test_device_cached_decoder_cuda_graphs_replay_across_batches():
    """cuda_graphs=True: the decoder call of every prefix length is captured once and replayed for the next batch
    of the same shape (other encoder outputs, other lengths, a primer of three tokens); results must equal the
    eager device-cache path."""
    from llm_guided_asr_b200 import BatchBeamSearch, CTCPrefixScorer
    from llm_guided_asr_b200.testing import synthetic as syn
    from llm_guided_asr_b200.device_cache import DeviceCachedDecoderScorer

    V, beam = 96, 3
    eos = V - 1
    # IdentityCTC reads the encoder output as (T, V) log-posteriors, so the decoder's memory width is V here
    dec = syn.TinyDecoder(V, D=V, seed=6).eval().to(DEV)

    def search(sc):
        bs = BatchBeamSearch(scorers=dict(decoder=sc, ctc=CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos)),
                             weights=dict(decoder=0.6, ctc=0.4), beam_size=beam, vocab_size=V, sos=eos, eos=eos,
                             pre_beam_score_key="full")
        bs.set_hyp_primer([eos, 7, 8])
        return bs

    graphed, eager = DeviceCachedDecoderScorer(dec, cuda_graphs=True), DeviceCachedDecoderScorer(dec)
    bs_g, bs_e = search(graphed), search(eager)
    for batch in range(3):
        lens = [30, 22 + batch]
        xs = [syn.random_logp(100 + 10 * batch + i, t, V, scale=2.0).to(DEV) for i, t in enumerate(lens)]
        a = bs_g.decode_batch(xs, -9, 0.0)
        b = bs_e.decode_batch(xs, -9, 0.0)
        for u, v in zip(a, b):
            assert len(u) == len(v) >= 1
            for h, r in zip(u, v):
                assert h.yseq.tolist() == r.yseq.tolist()
                assert abs(float(h.score) - float(r.score)) <= SCORE_ATOL
    assert len(graphed._graphs) >= 5  # captured once, replayed by the later batches


@pytest.mark.parametrize("graphs", [False, True])
def test_device_cached_decoder_and_lm_together(graphs):
    """Decoder and language model both behind the device protocol (two cached scorers, two fused terms) against
    the same search with the reference's host-side state protocol; two batches, the second replays graphs."""
    from llm_guided_asr_b200 import BatchBeamSearch, CTCPrefixScorer
    from llm_guided_asr_b200.testing import synthetic as syn
    from llm_guided_asr_b200.device_cache import DeviceCachedDecoderScorer, DeviceCachedLMScorer

    V, beam = 80, 4
    eos = V - 1
    dec = syn.TinyDecoder(V, D=V, seed=8).eval().to(DEV)
    lm = syn.TinyLM(V, seed=9).eval().to(DEV)
    weights = dict(decoder=0.6, ctc=0.4, lm=0.5)

    def search(d, m, **kw):
        return BatchBeamSearch(scorers=dict(decoder=d, ctc=CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos), lm=m),
                               weights=weights, beam_size=beam, vocab_size=V, sos=eos, eos=eos, pre_beam_score_key="full", **kw)

    dev_bs = search(DeviceCachedDecoderScorer(dec, cuda_graphs=graphs), DeviceCachedLMScorer(lm, cuda_graphs=graphs))
    host_bs = search(dec, lm, device_caches=False)
    assert isinstance(search(dec, lm).full_scorers["lm"], DeviceCachedLMScorer)  # bare TransformerLM surface: wrapped too
    for batch in range(2):
        lens = [26, 33, 19]
        xs = [syn.random_logp(200 + 10 * batch + i, t, V, scale=2.0).to(DEV) for i, t in enumerate(lens)]
        a = dev_bs.decode_batch(xs, -10, 0.0)
        b = host_bs.decode_batch(xs, -10, 0.0)
        for u, v in zip(a, b):
            assert len(u) == len(v) >= 1
            for h, r in zip(u, v):
                assert h.yseq.tolist() == r.yseq.tolist()
                assert abs(float(h.score) - float(r.score)) <= SCORE_ATOL
                for k in ("decoder", "lm", "ctc"):
                    assert abs(float(h.scores[k]) - float(r.scores[k])) <= SCORE_ATOL


def test_score_filter_matches_reference_goldens_and_oracle():
    import os

    from llm_guided_asr_b200.score_filter import ScoreFilter
    from oracle.score_filter_oracle import score_filter

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "score_filter.npz"))
    V, sos, eos, notime, first, last = [int(v) for v in g["spec"]]
    f = ScoreFilter(notime, first, last, sos, eos, V).to(DEV)
    for n in sorted(set(int(v) for v in g["lens"])):
        idx = [i for i in range(len(g["ys"])) if g["lens"][i] == n]
        ys = torch.as_tensor(g["ys"][idx, :n]).to(DEV)
        out, st = f.batch_score(ys, [None] * len(idx), None)
        out = out.cpu().numpy()
        assert np.all((out == 0) | np.isneginf(out)) and st == [None] * len(idx)
        assert np.array_equal(np.isneginf(out), g["masks"][idx])  # recorded from the reference class
        one, _ = f.score(ys[0], None, None)
        assert np.array_equal(np.isneginf(one.cpu().numpy()), g["masks"][idx[0]])
    # a vocabulary that is not a multiple of 4 and a strided prefix array, against the oracle
    V2 = 1003
    f2 = ScoreFilter(3, 500, 999, 1, 2, V2).to(DEV)
    rng = np.random.RandomState(4)
    big = rng.randint(4, V2, size=(6, 5, 20)).astype(np.int64)
    big[:, :, 0], big[:, :, 1], big[:, :, 2] = 1, 7, 8
    big[0, :, 2] = 3
    d_big = torch.as_tensor(big).to(DEV)
    for L in (3, 4, 9, 20):
        got = f2.batch_score_padded(d_big[:, :, :L], None, None).cpu().numpy()
        want = score_filter(big[:, :, :L].reshape(30, L), V2, 1, 2, 3, 500, 999)
        assert np.array_equal(got, want)


def test_fused_search_with_score_filter_and_primer():
    """OWSM-style decode: <sos><lang><task> primer, timestamp rules as an extra full scorer (weight 1.0) whose
    -inf entries go through the fusion, the pre-beam and the final top-k; against the oracle search."""
    from llm_guided_asr_b200 import BatchBeamSearch, CTCPrefixScorer
    from llm_guided_asr_b200.testing import synthetic as syn
    from llm_guided_asr_b200.score_filter import ScoreFilter
    from oracle.score_filter_oracle import score_filter

    V, beam, T = 200, 5, 60
    sos, eos, notime, first, last = 1, 2, 3, 100, 199
    logp = syn.random_logp(9, T, V, scale=2.0)
    table = syn.random_table(10, 64, 4, V, scale=1.5)
    weights = dict(decoder=0.7, ctc=0.3, scorefilter=1.0)
    primer = [sos, 5, 6]
    for prim in (primer, [sos, 5, notime]):
        bs = BatchBeamSearch(
            scorers=dict(decoder=syn.PrefixTableScorer(table.to(DEV)), ctc=CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos),
                         scorefilter=ScoreFilter(notime, first, last, sos, eos, V).to(DEV)),
            weights=weights, beam_size=beam, vocab_size=V, sos=sos, eos=eos, pre_beam_score_key="full")
        bs.set_hyp_primer(prim)
        got = bs(logp.to(DEV), -12, 0.0)
        from oracle.beam_search_oracle import torch_scorer_fn

        full = dict(decoder=torch_scorer_fn(syn.PrefixTableScorer(table)),
                    scorefilter=lambda ys: score_filter(np.asarray(ys), V, sos, eos, notime, first, last))
        orc = BeamSearchOracle(full, weights, beam, V, sos, eos, pre_beam_score_key="full", hyp_primer=prim)
        ref = orc.forward(logp.numpy(), -12, 0.0)
        assert len(got) == len(ref) >= 1
        for h, r in zip(got, ref):
            assert h.yseq.tolist() == r.yseq
            assert abs(float(h.score) - float(r.score)) <= SCORE_ATOL
