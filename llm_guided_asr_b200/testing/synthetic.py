"""Deterministic synthetic inputs and table-driven full scorers for parity tests and the bench.

There is no network for datasets or checkpoints, so every test / bench input is synthetic
(BASELINE.md §3).  The neural full scorers of the reference (attention decoder, LM, LLM;
SURVEY.md §8 a18) are *producers* of ``(n, V)`` log-prob matrices that stay in PyTorch; for
parity work they are replaced by table look-ups whose output bits are identical on every
device, so that the reference search (CPU), the oracle (CPU) and the CUDA path see exactly
the same ``(n, V)`` matrices.

All scorers here implement ``BatchScorerInterface`` (``espnet/nets/scorer_interface.py:83-126``)
and therefore plug into the unmodified reference ``BatchBeamSearch`` too.
"""

from typing import Any, List, Tuple

import numpy as np
import torch

from ..scorer_interface import BatchScorerInterface

LOGZERO = -10000000000.0  # espnet/nets/ctc_prefix_score.py:32


def prefix_hash(ys: torch.Tensor) -> torch.Tensor:
    """Integer hash of each int64 prefix row of ``ys (n, L)`` -> ``(n,)`` int64 (exact on any device)."""
    pos = torch.arange(1, ys.shape[1] + 1, device=ys.device, dtype=torch.int64)
    mixed = ((ys.to(torch.int64) + 1) * pos) % 1000003
    return (mixed * 40503).sum(dim=1) % 2147483647


class PrefixTableScorer(BatchScorerInterface):
    """Full scorer whose row is ``table[min(L-1, P-1), hash(prefix) % J]``.

    ``table`` is a ``(P, J, V)`` tensor of log-probabilities fixed in advance, so the
    scores of a prefix are bit-identical wherever they are computed.
    """

    def __init__(self, table: torch.Tensor):
        assert table.dim() == 3
        self.table = table

    def to(self, device):
        self.table = self.table.to(device)
        return self

    def score(self, y, state, x):
        s, _ = self.batch_score(y.unsqueeze(0), [state], None)
        return s[0], None

    def batch_score(self, ys: torch.Tensor, states: List[Any], xs: torch.Tensor) -> Tuple[torch.Tensor, List[Any]]:
        P, J, _ = self.table.shape
        p = min(ys.shape[1] - 1, P - 1)
        j = (prefix_hash(ys) % J).to(self.table.device)
        return self.table[p].index_select(0, j), [None] * ys.shape[0]


class SlotTableScorer(BatchScorerInterface):
    """Full scorer that returns ``tables[step, row0 : row0 + n]`` (positional; the bench workload).

    The matrix a hypothesis sees depends on the decode position and on its slot in the
    running beam, i.e. ``log_softmax(2*randn(B*W, V))`` with seed 11+step of BASELINE.md §3.
    """

    def __init__(self, tables: torch.Tensor, row0: int = 0, primer_len: int = 1):
        assert tables.dim() == 3
        self.tables = tables
        self.row0 = row0
        self.primer_len = primer_len

    def score(self, y, state, x):
        s, _ = self.batch_score(y.unsqueeze(0), [state], None)
        return s[0], None

    def batch_score(self, ys, states, xs):
        step = ys.shape[1] - self.primer_len
        n = ys.shape[0]
        return self.tables[step, self.row0 : self.row0 + n], [None] * n


class IdentityCTC(torch.nn.Module):
    """Stand-in for the CTC head (``espnet2/asr/ctc.py:196-204``): the "encoder output" already
    is the ``(T, V)`` log-posterior matrix, ``log_softmax`` returns a copy of it."""

    def log_softmax(self, hs_pad: torch.Tensor) -> torch.Tensor:
        return hs_pad.clone()

    def argmax(self, hs_pad: torch.Tensor) -> torch.Tensor:
        return torch.argmax(hs_pad, dim=2)


def random_logp(seed: int, T: int, V: int, scale: float = 3.0) -> torch.Tensor:
    """``log_softmax(scale * randn(T, V))`` in fp32 on the CPU (BASELINE.md §3 micro-benchmark input)."""
    g = torch.Generator().manual_seed(seed)
    return torch.log_softmax(scale * torch.randn(T, V, generator=g), dim=-1)


def random_table(seed: int, P: int, J: int, V: int, scale: float = 2.0) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    return torch.log_softmax(scale * torch.randn(P, J, V, generator=g), dim=-1)


def peaky_case(seed: int, T: int, V: int, ntok: int, eos: int, J: int = 4, sharp: float = 9.0):
    """A "trained-like" case: a hidden token sequence, CTC posteriors peaked along one
    alignment of it (blank elsewhere) and a full-scorer table peaked on the same sequence,
    then on ``eos``.  The search ends by ``end_detect`` after about ``ntok`` positions.

    Returns ``(logp (T, V), table (P, J, V), truth list[int])``.
    """
    rng = np.random.default_rng(seed)
    truth = [int(t) for t in rng.integers(1, V - 1, size=ntok)]
    truth = [t if t != eos else 1 for t in truth]
    # alignment: each token gets 1-2 frames, blanks fill the rest
    frames = np.zeros(T, dtype=np.int64)  # blank = 0
    slots = np.sort(rng.choice(np.arange(1, T - 1), size=ntok, replace=False))
    for s, tok in zip(slots, truth):
        frames[s] = tok
    g = torch.Generator().manual_seed(seed + 1)
    logits = torch.randn(T, V, generator=g)
    logits[torch.arange(T), torch.as_tensor(frames)] += sharp
    logp = torch.log_softmax(logits, dim=-1)
    P = ntok + 4
    tl = 1.5 * torch.randn(P, J, V, generator=g)
    for p in range(P):
        tl[p, :, truth[p] if p < ntok else eos] += sharp
    table = torch.log_softmax(tl, dim=-1)
    return logp, table, truth


__all__ = [
    "LOGZERO",
    "prefix_hash",
    "PrefixTableScorer",
    "SlotTableScorer",
    "IdentityCTC",
    "random_logp",
    "random_table",
    "peaky_case",
]


class _TinyDecoderLayer(torch.nn.Module):
    """Pre-norm decoder layer with the call contract of the reference's ``DecoderLayer``
    (``espnet/nets/pytorch_backend/transformer/decoder_layer.py``): ``layer(tgt, tgt_mask, memory,
    memory_mask, cache=None) -> (x, tgt_mask, memory, memory_mask)``; with ``cache (n, L-1, D)`` only the
    last position is computed and the result is ``cat([cache, x_last], dim=1)``."""

    def __init__(self, D: int, heads: int, ff: int):
        super().__init__()
        self.H = heads
        self.n1, self.n2, self.n3 = torch.nn.LayerNorm(D), torch.nn.LayerNorm(D), torch.nn.LayerNorm(D)
        self.q1, self.kv1, self.o1 = torch.nn.Linear(D, D), torch.nn.Linear(D, 2 * D), torch.nn.Linear(D, D)
        self.q2, self.kv2, self.o2 = torch.nn.Linear(D, D), torch.nn.Linear(D, 2 * D), torch.nn.Linear(D, D)
        self.f1, self.f2 = torch.nn.Linear(D, ff), torch.nn.Linear(ff, D)

    def _att(self, q, kv, mask):
        n, Lq, D = q.shape
        k, v = kv.chunk(2, dim=-1)
        sp = lambda t: t.view(n, t.shape[1], self.H, D // self.H).transpose(1, 2)  # noqa: E731
        a = torch.matmul(sp(q), sp(k).transpose(-1, -2)) / float(D // self.H) ** 0.5
        if mask is not None:
            a = a.masked_fill(~mask.unsqueeze(1), float("-inf"))
        return torch.matmul(torch.softmax(a, dim=-1), sp(v)).transpose(1, 2).reshape(n, Lq, D)

    def forward(self, tgt, tgt_mask, memory, memory_mask, cache=None):
        y = self.n1(tgt)
        if cache is None:
            q_in, res, m1 = y, tgt, tgt_mask
        else:
            q_in, res = y[:, -1:], tgt[:, -1:]
            m1 = None if tgt_mask is None else tgt_mask[:, -1:]
        x = res + self.o1(self._att(self.q1(q_in), self.kv1(y), m1))
        x = x + self.o2(self._att(self.q2(self.n2(x)), self.kv2(memory), memory_mask))
        x = x + self.f2(torch.relu(self.f1(self.n3(x))))
        if cache is not None:
            x = torch.cat([cache, x], dim=1)
        return x, tgt_mask, memory, memory_mask


class TinyDecoder(torch.nn.Module, BatchScorerInterface):
    """Small random-init attention decoder with the attribute and method surface of the reference's
    ``BaseTransformerDecoder`` (``espnet2/asr/decoder/transformer_decoder.py:36-289``: ``embed``, ``decoders``,
    ``after_norm``, ``normalize_before``, ``output_layer``, ``forward_one_step``, ``score``, ``batch_score`` with
    per-hypothesis lists of per-layer ``(L, D)`` states).  Stands in for it on a box without the reference."""

    def __init__(self, V: int, D: int = 32, layers: int = 2, heads: int = 4, ff: int = 64, seed: int = 0):
        torch.nn.Module.__init__(self)
        with torch.random.fork_rng(devices=[]):
            torch.manual_seed(seed)
            self.embed = torch.nn.Embedding(V, D)
            self.decoders = torch.nn.ModuleList([_TinyDecoderLayer(D, heads, ff) for _ in range(layers)])
            self.after_norm = torch.nn.LayerNorm(D)
            self.output_layer = torch.nn.Linear(D, V)
        self.normalize_before = True

    def forward_one_step(self, tgt, tgt_mask, memory, memory_mask=None, *, cache=None):
        x = self.embed(tgt)
        if cache is None:
            cache = [None] * len(self.decoders)
        new_cache = []
        for c, layer in zip(cache, self.decoders):
            x, tgt_mask, memory, memory_mask = layer(x, tgt_mask, memory, memory_mask, cache=c)
            new_cache.append(x)
        y = self.after_norm(x[:, -1]) if self.normalize_before else x[:, -1]
        return torch.log_softmax(self.output_layer(y), dim=-1), new_cache

    def score(self, ys, state, x):
        mask = torch.ones(ys.shape[0], ys.shape[0], dtype=torch.bool, device=ys.device).tril().unsqueeze(0)
        logp, st = self.forward_one_step(ys.unsqueeze(0), mask, x.unsqueeze(0), cache=state)
        return logp.squeeze(0), st

    def batch_score(self, ys, states, xs):
        n, L = ys.shape
        nl = len(self.decoders)
        cache = None if states[0] is None else [torch.stack([states[b][i] for b in range(n)]) for i in range(nl)]
        mask = torch.ones(L, L, dtype=torch.bool, device=ys.device).tril().unsqueeze(0)
        logp, st = self.forward_one_step(ys, mask, xs, cache=cache)
        return logp, [[st[i][b] for i in range(nl)] for b in range(n)]


class _TinyLMLayer(torch.nn.Module):
    """Pre-norm self-attention layer with the call contract of the reference's ``EncoderLayer``
    (``espnet/nets/pytorch_backend/transformer/encoder_layer.py``): ``layer(x, mask, cache=None) -> (x, mask)``;
    with ``cache (n, L-1, D)`` only the last position is computed and appended."""

    def __init__(self, D: int, heads: int, ff: int):
        super().__init__()
        self.H = heads
        self.n1, self.n2 = torch.nn.LayerNorm(D), torch.nn.LayerNorm(D)
        self.q, self.kv, self.o = torch.nn.Linear(D, D), torch.nn.Linear(D, 2 * D), torch.nn.Linear(D, D)
        self.f1, self.f2 = torch.nn.Linear(D, ff), torch.nn.Linear(ff, D)

    def forward(self, x, mask, cache=None):
        y = self.n1(x)
        if cache is None:
            q_in, res, m1 = y, x, mask
        else:
            q_in, res = y[:, -1:], x[:, -1:]
            m1 = None if mask is None else mask[:, -1:]
        n, Lq, D = q_in.shape
        k, v = self.kv(y).chunk(2, dim=-1)
        sp = lambda t: t.view(n, t.shape[1], self.H, D // self.H).transpose(1, 2)  # noqa: E731
        a = torch.matmul(sp(self.q(q_in)), sp(k).transpose(-1, -2)) / float(D // self.H) ** 0.5
        if m1 is not None:
            a = a.masked_fill(~m1.unsqueeze(1), float("-inf"))
        att = torch.matmul(torch.softmax(a, dim=-1), sp(v)).transpose(1, 2).reshape(n, Lq, D)
        h = res + self.o(att)
        h = h + self.f2(torch.relu(self.f1(self.n2(h))))
        if cache is not None:
            h = torch.cat([cache, h], dim=1)
        return h, mask


class _TinyLMEncoder(torch.nn.Module):
    def __init__(self, E: int, D: int, layers: int, heads: int, ff: int):
        super().__init__()
        self.embed = torch.nn.Linear(E, D)  # input_layer="linear" of the reference's Encoder
        self.encoders = torch.nn.ModuleList([_TinyLMLayer(D, heads, ff) for _ in range(layers)])
        self.after_norm = torch.nn.LayerNorm(D)
        self.normalize_before = True

    def forward_one_step(self, xs, masks, *, cache=None):
        xs = self.embed(xs)
        if cache is None:
            cache = [None] * len(self.encoders)
        new_cache = []
        for c, e in zip(cache, self.encoders):
            xs, masks = e(xs, masks, cache=c)
            new_cache.append(xs)
        return self.after_norm(xs), masks, new_cache


class TinyLM(torch.nn.Module, BatchScorerInterface):
    """Small random-init language model with the attribute and method surface of the reference's
    ``TransformerLM`` (``espnet2/lm/transformer_lm.py:12-133``: ``embed``, ``encoder``, ``decoder``, ``_target_mask``,
    ``score``, ``batch_score`` with per-hypothesis lists of per-layer states)."""

    def __init__(self, V: int, E: int = 16, D: int = 32, layers: int = 2, heads: int = 4, ff: int = 64, seed: int = 0):
        torch.nn.Module.__init__(self)
        with torch.random.fork_rng(devices=[]):
            torch.manual_seed(seed)
            self.embed = torch.nn.Embedding(V, E)
            self.encoder = _TinyLMEncoder(E, D, layers, heads, ff)
            self.decoder = torch.nn.Linear(D, V)

    def _target_mask(self, ys):
        L = ys.shape[-1]
        return (ys != 0).unsqueeze(-2) & torch.ones(L, L, dtype=torch.bool, device=ys.device).tril().unsqueeze(0)

    def score(self, y, state, x):
        y = y.unsqueeze(0)
        h, _, cache = self.encoder.forward_one_step(self.embed(y), self._target_mask(y), cache=state)
        return self.decoder(h[:, -1]).log_softmax(dim=-1).squeeze(0), cache

    def batch_score(self, ys, states, xs):
        n, nl = len(ys), len(self.encoder.encoders)
        cache = None if states[0] is None else [torch.stack([states[b][i] for b in range(n)]) for i in range(nl)]
        h, _, st = self.encoder.forward_one_step(self.embed(ys), self._target_mask(ys), cache=cache)
        logp = self.decoder(h[:, -1]).log_softmax(dim=-1)
        return logp, [[st[i][b] for i in range(nl)] for b in range(n)]
