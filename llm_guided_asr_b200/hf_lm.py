"""Hugging Face causal language models as beam-search full scorers (the shallow-fusion LM / LLM of configs[2], [4]).

``HFPrefixLM`` restates the reference's ``HuggingfaceOPTModel`` scorer surface
(``espnet2/lm/huggingface_pretrained_opt_lm.py:20-148``: a ``transformers`` base model in ``self.decoder``, a bias-free
``self.lm_head``, ``score`` / ``batch_score`` that re-read the whole prefix and return the last position's
``log_softmax``) around ANY ``transformers`` base model, so that OPT, GPT-2 and Llama-style models plug into either
search through the plain ``BatchScorerInterface``; ``device_cache.DevicePrefixLMScorer`` wraps it for the device protocol
(one call per position for all utterances, logits handed to the fused step).

The reference ships no GPT-2 wrapper: ``gpt2_small(vocab)`` builds configs[2]'s scorer random-init
(no download), ``GPT2Config(vocab_size=V, n_layer=12, n_embd=768, n_head=12)``.
"""

from typing import Any, List, Tuple

import torch

from .scorer_interface import BatchScorerInterface


class HFPrefixLM(torch.nn.Module, BatchScorerInterface):
    """``decoder``: a ``transformers`` base model (``OPTModel``, ``GPT2Model``, ``LlamaModel`` ...); ``lm_head``:
    ``Linear(hidden, vocab, bias=False)`` as in the reference (``huggingface_pretrained_opt_lm.py:43-45``)."""

    def __init__(self, decoder: torch.nn.Module, hidden_size: int, vocab_size: int):
        torch.nn.Module.__init__(self)
        self.decoder = decoder
        self.lm_head = torch.nn.Linear(hidden_size, vocab_size, bias=False)

    def forward(self, input: torch.Tensor, hidden: None = None):  # huggingface_pretrained_opt_lm.py:52-70
        y = self.decoder(input, attention_mask=input != 0, return_dict=True).last_hidden_state
        return self.lm_head(y), None

    def score(self, y: torch.Tensor, state: Any, x: torch.Tensor) -> Tuple[torch.Tensor, Any]:
        logp, _ = self.batch_score(y.unsqueeze(0), [state], None)
        return logp.squeeze(0), None

    def batch_score(self, ys: torch.Tensor, states: List[Any], xs: torch.Tensor) -> Tuple[torch.Tensor, List[Any]]:
        """``ys (n, L)`` int64 prefixes -> ``(n, V)`` log-probabilities of the next token; no state is kept
        (``huggingface_pretrained_opt_lm.py:110-148``)."""
        h = self.decoder(ys, use_cache=False, return_dict=True).last_hidden_state
        return self.lm_head(h[:, -1]).log_softmax(dim=-1), [None] * len(ys)


def gpt2_small(vocab_size: int, n_positions: int = 1024, seed: int = 0, dtype=torch.float32, **overrides) -> HFPrefixLM:
    """GPT-2-small (12 layers, 768 wide, 12 heads) with a ``vocab_size`` embedding, random-init (there is no network for
    checkpoints): the LM / LLM shallow-fusion scorer of BASELINE.json configs[2] and [4]."""
    from transformers import GPT2Config, GPT2Model

    cfg = dict(vocab_size=vocab_size, n_positions=n_positions, n_embd=768, n_layer=12, n_head=12, resid_pdrop=0.0,
               embd_pdrop=0.0, attn_pdrop=0.0)
    cfg.update(overrides)
    with torch.random.fork_rng(devices=[]):
        torch.manual_seed(seed)
        lm = HFPrefixLM(GPT2Model(GPT2Config(**cfg)), cfg["n_embd"], vocab_size)
    return lm.to(dtype).eval()
