"""CPU: the C-ABI library loads and exports every symbol include/lgasr.h declares; host-side logic
(length bounds, scorer partition, error behaviour) mirrors the reference; the product path fails
loudly without a CUDA device (no CPU fallback)."""

import ctypes as C
import os
import re

import pytest
import torch

from llm_guided_asr_b200 import BatchBeamSearch, CTCPrefixScorer, LengthBonus, _lib
from llm_guided_asr_b200.testing import synthetic as syn
from llm_guided_asr_b200.beam_search import length_bounds

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "lgasr.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lgasr_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"liblgasr.so does not export {n}"
    assert lib.lgasr_version() >= 100


def test_no_torch_types_in_header():
    src = open(os.path.join(ROOT, "include", "lgasr.h")).read()
    assert "at::" not in src and "torch::" not in src and "#include <torch" not in src


def test_create_fails_without_cuda_device():
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    lib = _lib.load()
    cfg = _lib.Config(n_utts=1, max_frames=8, vocab=10, beam=2, pre_beam=3, n_terms=1, pre_beam_key=-1, use_ctc=1,
                      ctc_weight=0.3, blank=0, sos=9, eos=9, max_len_cap=8, primer_len=1, end_detect=1)
    h = C.c_void_p()
    rc = lib.lgasr_create(C.byref(cfg), C.byref(h))
    assert rc == -2  # LGASR_ECUDA: no CPU fallback
    with pytest.raises(_lib.LgasrError):
        _lib.check(rc)


def test_create_rejects_bad_config():
    lib = _lib.load()
    cfg = _lib.Config(n_utts=0, max_frames=8, vocab=10, beam=2, pre_beam=3, n_terms=1, pre_beam_key=-1, use_ctc=1,
                      ctc_weight=0.3, blank=0, sos=9, eos=9, max_len_cap=8, primer_len=1, end_detect=1)
    h = C.c_void_p()
    assert lib.lgasr_create(C.byref(cfg), C.byref(h)) == -1  # LGASR_EINVAL before any device work


def test_length_bounds_follow_beam_search_forward():
    # espnet/nets/beam_search.py:417-427
    assert length_bounds(100, 0.0, 0.0) == (100, 0)
    assert length_bounds(100, -64, 0.0) == (64, 0)
    assert length_bounds(100, 0.5, 0.1) == (50, 10)
    assert length_bounds(1, 0.3, 0.0) == (1, 0)
    assert length_bounds(100, 0.0, -5)[1] == 5


def _scorers(V):
    table = syn.random_table(3, 8, 2, V)
    return dict(decoder=syn.PrefixTableScorer(table), ctc=CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=V - 1),
                length_bonus=LengthBonus(V))


def test_scorer_partition_and_errors_mirror_reference():
    V = 20
    sc = _scorers(V)
    bs = BatchBeamSearch(scorers=sc, weights=dict(decoder=0.7, ctc=0.3, length_bonus=0.0), beam_size=4, vocab_size=V,
                         sos=V - 1, eos=V - 1, pre_beam_score_key="full")
    assert list(bs.full_scorers) == ["decoder"]  # weight 0 dropped entirely (beam_search.py:82-85)
    assert list(bs.part_scorers) == ["ctc"]
    assert bs.pre_beam_size == 6 and bs.do_pre_beam
    with pytest.raises(KeyError):
        BatchBeamSearch(scorers=sc, weights=dict(decoder=0.7, ctc=0.3), beam_size=4, vocab_size=V, sos=0, eos=0,
                        pre_beam_score_key="nope")
    with pytest.raises(AssertionError):
        BatchBeamSearch(scorers=dict(decoder=object()), weights=dict(decoder=1.0), beam_size=4, vocab_size=V, sos=0, eos=0)
    # CPU tensors are refused: there is no CPU fallback
    with pytest.raises(RuntimeError):
        bs(torch.zeros(5, V), 0.0, 0.0)


def test_state_side_entry_points_validate_before_any_device_work():
    # include/lgasr.h: lgasr_log_softmax_rows / lgasr_reorder_rows refuse bad arguments with LGASR_EINVAL
    lib = _lib.load()
    assert lib.lgasr_log_softmax_rows(None, 0, 10, None, 10, 1, 10, None) == -1
    assert lib.lgasr_reorder_rows(None, None, None, 1, 0, 0, 0, 0, 0, None) == -1
    from llm_guided_asr_b200.device_cache import DeviceCachedDecoderScorer, log_softmax_rows

    with pytest.raises(RuntimeError):
        log_softmax_rows(torch.zeros(2, 8))  # CPU tensor: no fallback
    with pytest.raises(TypeError):
        DeviceCachedDecoderScorer(torch.nn.Linear(2, 2))
    # the wrapper delegates the plain scorer protocol to the decoder it wraps
    dec = syn.TinyDecoder(17, D=16, layers=1, heads=2, ff=16)
    wrap = DeviceCachedDecoderScorer(dec)
    ys = torch.tensor([[16, 3], [16, 5]])
    mem = torch.randn(2, 6, 16)
    a, _ = wrap.batch_score(ys, [None, None], mem)
    b, _ = dec.batch_score(ys, [None, None], mem)
    assert torch.equal(a, b)
