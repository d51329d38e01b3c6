"""CPU, build container only (skipped where /root/reference is absent): the network bodies the device-protocol
wrappers run (`_net`) against the REFERENCE's own modules -- `TransformerDecoder.forward_one_step`
(espnet2/asr/decoder/transformer_decoder.py:169-217) and `TransformerLM.batch_score`
(espnet2/lm/transformer_lm.py:95-133) -- with and without caches.  Only the arithmetic-free plumbing of the
wrappers is exercised here; the CUDA kernels they call are covered by the GPU tests."""

import os
import sys

import pytest
import torch

REF = "/root/reference"
if not os.path.isdir(REF):
    pytest.skip("reference tree not present on this machine", allow_module_level=True)
sys.path.insert(0, REF)
sys.dont_write_bytecode = True
try:
    from espnet2.asr.decoder.transformer_decoder import TransformerDecoder
    from espnet2.lm.transformer_lm import TransformerLM
except Exception as e:  # pragma: no cover
    pytest.skip(f"reference modules do not import here: {e}", allow_module_level=True)

from llm_guided_asr_b200.device_cache import DeviceCachedDecoderScorer, DeviceCachedLMScorer


def test_decoder_body_matches_forward_one_step():
    torch.manual_seed(0)
    V, D, n, T = 50, 32, 5, 11
    dec = TransformerDecoder(V, D, attention_heads=4, linear_units=64, num_blocks=2, dropout_rate=0.0,
                             positional_dropout_rate=0.0, self_attention_dropout_rate=0.0,
                             src_attention_dropout_rate=0.0).eval()
    w = DeviceCachedDecoderScorer(dec, memory_kv=False)
    memory = torch.randn(n, T, D)
    w._memory, w._memory_mask, w._fast = memory, torch.ones(n, 1, T, dtype=torch.bool), False
    ys = torch.randint(1, V, (n, 4))
    with torch.no_grad():
        cache = [None] * 2
        for L in range(1, 5):
            tgt = ys[:, :L]
            mask = torch.ones(L, L, dtype=torch.bool).tril().unsqueeze(0)
            want, want_cache = dec.forward_one_step(tgt, mask, memory, cache=None if L == 1 else cache)
            logits, new = w._net(tgt, cache if L > 1 else [None] * 2)
            assert torch.allclose(torch.log_softmax(logits, dim=-1), want, atol=1e-5)
            for a, b in zip(new, want_cache):
                assert torch.allclose(a, b, atol=1e-5)
            cache = new


def test_decoder_memory_kv_path_matches_forward_one_step():
    """memory_kv: source-attention keys / values projected once per utterance; rows of two utterances of different
    lengths against the reference's forward_one_step on the padded, masked memory."""
    torch.manual_seed(3)
    V, D, B, W, T = 50, 32, 2, 3, 11
    lens = [11, 8]
    dec = TransformerDecoder(V, D, attention_heads=4, linear_units=64, num_blocks=2, dropout_rate=0.0,
                             positional_dropout_rate=0.0, self_attention_dropout_rate=0.0,
                             src_attention_dropout_rate=0.0).eval()
    xs = [torch.randn(t, D) for t in lens]
    w = DeviceCachedDecoderScorer(dec)
    n = B * W
    memory = torch.zeros(n, T, D)
    mmask = torch.zeros(n, 1, T, dtype=torch.bool)
    for b in range(B):
        memory[b * W : (b + 1) * W, : lens[b]] = xs[b]
        mmask[b * W : (b + 1) * W, :, : lens[b]] = True
    ys = torch.randint(1, V, (n, 4))
    with torch.no_grad():
        w._begin_inputs(xs, lens, B, W, T, torch.device("cpu"), True)
        assert w._fast
        cache, ref_cache = [None] * 2, None
        for L in range(1, 5):
            tgt = ys[:, :L]
            mask = torch.ones(L, L, dtype=torch.bool).tril().unsqueeze(0)
            want, ref_cache = dec.forward_one_step(tgt, mask, memory, mmask, cache=ref_cache)
            logits, cache = w._net(tgt, cache)  # the fast path caches self-attention keys / values, not layer outputs
            assert torch.allclose(torch.log_softmax(logits, dim=-1), want, atol=1e-5)
            assert cache[0].shape == (n, L, 2 * D)
        # a fresh batch that starts from a longer prefix (no cache yet): all positions are computed
        logits, new = w._net(ys, [None] * 2)
        want, _ = dec.forward_one_step(ys, torch.ones(4, 4, dtype=torch.bool).tril().unsqueeze(0), memory, mmask, cache=None)
        assert torch.allclose(torch.log_softmax(logits, dim=-1), want, atol=1e-5)


def test_lm_body_matches_batch_score():
    torch.manual_seed(1)
    V, n = 40, 4
    lm = TransformerLM(V, embed_unit=16, att_unit=32, head=2, unit=48, layer=2, dropout_rate=0.0,
                       positional_dropout_rate=0.0, attention_dropout_rate=0.0).eval()
    w = DeviceCachedLMScorer(lm)
    ys = torch.randint(1, V, (n, 5))
    with torch.no_grad():
        states = [None] * n
        cache = [None] * 2
        for L in range(1, 6):
            want, states = lm.batch_score(ys[:, :L], states, None)
            logits, cache = w._net(ys[:, :L], cache if L > 1 else [None] * 2)
            assert torch.allclose(torch.log_softmax(logits, dim=-1), want, atol=1e-5)
