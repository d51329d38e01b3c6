"""Full-scorer state kept on the device, next to the fused step.

What the reference does around every decoder call, per decode position:

* ``BaseTransformerDecoder.batch_score`` stacks the per-hypothesis, per-layer caches into batch tensors
  (``espnet2/asr/decoder/transformer_decoder.py:268-272``), calls ``forward_one_step`` (``:169-217``, ending in
  ``torch.log_softmax`` ``:212``) and un-stacks the new caches into Python lists again (``:285-286``);
* ``BatchBeamSearch.search`` then picks each survivor's state with ``select_state`` from Python
  (``espnet/nets/batch_beam_search.py:347-355``), after the parents have been copied to the host.

Here the caches of all ``B*beam`` row slots live in two persistent device buffers per scorer and are
re-ordered by ONE parent-indexed gather per layer (``lgasr_reorder_rows``) that reads the parents where
``k_step`` wrote them; the logits go through ``lgasr_log_softmax_rows`` straight into the row layout
``lgasr_step`` consumes.  The decoder network itself stays in torch (§8 a18).
"""

import ctypes as C
from typing import Any, List, Optional, Sequence, Tuple

import torch

from . import _lib
from .scorer_interface import BatchScorerInterface


def log_softmax_rows(logits: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``log_softmax(logits, dim=-1)`` in fp32 for a ``(rows, V)`` fp32 or bf16 CUDA tensor (rows may be
    strided), by ``lgasr_log_softmax_rows``.  ``out``: optional ``(rows, V)`` fp32 destination."""
    if logits.device.type != "cuda":
        raise RuntimeError("log_softmax_rows needs a CUDA tensor: there is no CPU fallback")
    if logits.dim() != 2 or logits.stride(1) != 1:
        raise ValueError("log_softmax_rows: (rows, V) tensor with unit stride along V expected")
    if logits.dtype == torch.float32:
        dt = 0
    elif logits.dtype == torch.bfloat16:
        dt = 1
    else:
        raise ValueError("log_softmax_rows: fp32 or bf16 logits expected")
    rows, V = logits.shape
    if out is None:
        out = torch.empty(rows, V, dtype=torch.float32, device=logits.device)
    if out.shape != (rows, V) or out.dtype != torch.float32 or out.stride(1) != 1 or out.device != logits.device:
        raise ValueError("log_softmax_rows: out must be a (rows, V) fp32 tensor on the same device")
    lib = _lib.load()
    with torch.cuda.device(logits.device):
        st = int(torch.cuda.current_stream(logits.device).cuda_stream)
        ld_in = logits.stride(0) if rows > 1 else V
        ld_out = out.stride(0) if rows > 1 else V
        _lib.check(lib.lgasr_log_softmax_rows(C.c_void_p(logits.data_ptr()), dt, ld_in, C.c_void_p(out.data_ptr()), ld_out,
                                              rows, V, st))
    return out


def row_logsumexp(logits: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``logsumexp(logits, dim=-1)`` in fp32 for a ``(rows, V)`` fp32 or bf16 CUDA tensor, by ``lgasr_row_logsumexp``: the
    per-row normaliser ``lgasr_step_ex`` subtracts while it reads the logits themselves (no ``(rows, V)`` fp32
    log-probability matrix is ever written)."""
    if logits.device.type != "cuda":
        raise RuntimeError("row_logsumexp needs a CUDA tensor: there is no CPU fallback")
    if logits.dim() != 2 or logits.stride(1) != 1 or logits.dtype not in (torch.float32, torch.bfloat16):
        raise ValueError("row_logsumexp: (rows, V) fp32 / bf16 tensor with unit stride along V expected")
    rows, V = logits.shape
    if out is None:
        out = torch.empty(rows, dtype=torch.float32, device=logits.device)
    lib = _lib.load()
    with torch.cuda.device(logits.device):
        st = int(torch.cuda.current_stream(logits.device).cuda_stream)
        ld_in = logits.stride(0) if rows > 1 else V
        _lib.check(lib.lgasr_row_logsumexp(C.c_void_p(logits.data_ptr()), 1 if logits.dtype == torch.bfloat16 else 0, ld_in,
                                           C.c_void_p(out.data_ptr()), rows, V, st))
    return out


class _InPlaceCaches(list):
    """Per-layer caches that are views ``[:, :L]`` of the scorer's current persistent buffer (appended in place)."""


class _DeviceCachedScorer(torch.nn.Module, BatchScorerInterface):
    """Machinery shared by the device-protocol scorers: static buffers per batch shape, one network call per
    position for all ``B*beam`` row slots (``_net``), ``lgasr_log_softmax_rows`` into the row layout of
    ``lgasr_step``, ``lgasr_reorder_rows`` of every layer's cache by the parents the step chose, optional CUDA-graph
    capture of the network call per prefix length."""

    def __init__(self, net: torch.nn.Module, fused_log_softmax: bool = True, cuda_graphs: bool = False,
                 fused_normaliser: bool = False):
        torch.nn.Module.__init__(self)
        self.net = net
        self.fused_log_softmax = fused_log_softmax
        # fused_normaliser: hand the LOGITS (fp32 or bf16, as the network produced them) plus one logsumexp per row to
        # the step (lgasr_row_logsumexp + lgasr_step_ex) instead of writing an fp32 (B*W, V) log-probability matrix:
        # the step reads the row once, in its own type (2 bytes per entry for a bf16 LLM head)
        self.fused_normaliser = fused_normaliser
        # cuda_graphs: the network call of a position (embed -> layers -> output layer -> log-softmax) is captured
        # once per prefix length into a CUDA graph over static buffers and replayed for every later batch of the
        # same shape: an eager step is ~25 launches per layer, i.e. launch-bound at beam-search sizes
        self.cuda_graphs = cuda_graphs
        self._eng = None
        self._shape = None
        self._graphs = {}
        self._pool = None
        self._slots, self._inplace = None, False

    # ---- plain scorer protocol: delegate to the wrapped network --------------------------------------------
    def batch_init_state(self, x):
        return self.net.batch_init_state(x) if hasattr(self.net, "batch_init_state") else None

    def select_state(self, state, i, new_id=None):
        if state is None:
            return None
        return self.net.select_state(state, i) if hasattr(self.net, "select_state") else state[i]

    def score(self, ys, state, x):
        return self.net.score(ys, state, x)

    def batch_score(self, ys: torch.Tensor, states: List[Any], xs: torch.Tensor) -> Tuple[torch.Tensor, List[Any]]:
        return self.net.batch_score(ys, states, xs)

    # ---- hooks ----------------------------------------------------------------------------------------------
    def _n_layers(self) -> int:
        raise NotImplementedError

    def _begin_inputs(self, xs, lens, B, W, T, dev, fresh_shape: bool) -> None:
        """Per-batch inputs of the network other than the prefixes (static buffers)."""

    def _net(self, tgt: torch.Tensor, cache: List[Optional[torch.Tensor]]):
        """``tgt (B*W, L)`` int64, per-layer caches ``(B*W, L-1, Dm)`` or None -> ``(logits (B*W, V), new caches)``."""
        raise NotImplementedError

    # ---- device protocol ------------------------------------------------------------------------------------
    def device_begin(self, xs: Sequence[torch.Tensor], lens: Sequence[int], engine, max_positions: int) -> None:
        """Start a batch: ``xs[b] (T_b, D)`` encoder outputs; row slot ``b*beam + j`` belongs to utterance ``b``."""
        B, W = engine.cfg.n_utts, engine.cfg.beam
        dev = xs[0].device
        T = max(int(t) for t in lens)
        shape = (B, W, T, int(xs[0].shape[-1]), int(max_positions), engine.cfg.vocab, xs[0].dtype, dev)
        fresh_shape = shape != self._shape
        if fresh_shape:  # static buffers (and the graphs captured over them) live as long as the shape
            self._shape = shape
            self._graphs, self._pool = {}, None
            self._tgt = torch.zeros(B * W, int(max_positions), dtype=torch.int64, device=dev)
            self._out = None if self.fused_normaliser and engine.cfg.vocab % 4 == 0 else torch.empty(
                B * W, engine.cfg.vocab, dtype=torch.float32, device=dev)
            self._lse = torch.empty(B * W, dtype=torch.float32, device=dev)
            self._buf = None  # [2] x (n_layers, B*W, cap, Dm), allocated on the first call (Dm known then)
        self._begin_inputs(xs, lens, B, W, T, dev, fresh_shape)
        self._eng = engine
        self._cap = int(max_positions)
        self._pp = 0
        self._new = None
        self._fresh = True  # no cache yet in this batch (the persistent buffers may hold the previous batch)
        self._V = engine.cfg.vocab

    def _forward(self, L: int):
        """Network call for prefix length ``L`` over the static buffers; writes ``self._out``, returns the new
        per-layer caches ``(B*W, L, Dm)``."""
        tgt = self._tgt[:, :L]
        nl = self._n_layers()
        self._slots = None
        if self._fresh or self._buf is None or L == 1:
            cache = [None] * nl
        else:
            cache = [self._buf[self._pp][i, :, : L - 1] for i in range(nl)]
            # where a network can append its new cache entry in place (position L - 1 of the persistent buffer it reads
            # the earlier positions from) instead of returning a concatenated copy
            self._slots = [self._buf[self._pp][i, :, :L] for i in range(nl)]
        self._inplace = False
        logits, new = self._net(tgt, cache)
        if self._inplace:
            new = _InPlaceCaches(new)
        if logits.dtype not in (torch.float32, torch.bfloat16):
            logits = logits.float()
        if self._out is None:  # fused normaliser: (logits, logsumexp) go to the step as they are
            logits = logits.contiguous()
            row_logsumexp(logits, self._lse)
            self._ret = (logits, self._lse)
        elif self.fused_log_softmax:
            log_softmax_rows(logits, self._out)
            self._ret = self._out
        else:
            self._out.copy_(torch.log_softmax(logits.float(), dim=-1))
            self._ret = self._out
        return new

    def batch_score_padded(self, yseq: torch.Tensor, n_run, done) -> torch.Tensor:
        """``yseq (B, W, L)`` int64 running prefixes (all row slots) -> ``(B*W, V)`` fp32 log-probabilities, or (with
        ``fused_normaliser``) the pair ``(logits (B*W, V) fp32 / bf16, logsumexp (B*W,) fp32)``; rows of slots that
        are not running hold values nobody reads."""
        B, W, L = yseq.shape
        self._tgt[:, :L].copy_(yseq.reshape(B * W, L).clamp(0, self._V - 1))  # idle slots may hold stale ids
        stateless = self._n_layers() == 0
        if not self.cuda_graphs or self._fresh or (self._buf is None and not stateless):
            self._new = self._forward(L)
            if stateless:
                self._fresh = False
            return self._ret
        key = (L, self._pp)
        ent = self._graphs.get(key)
        if ent is None:
            cur = torch.cuda.current_stream()
            side = torch.cuda.Stream()
            side.wait_stream(cur)
            with torch.cuda.stream(side):  # warm-up outside the capture (library workspaces, lazy kernels)
                self._forward(L)
            cur.wait_stream(side)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, pool=self._pool):
                new = self._forward(L)
            if self._pool is None:
                self._pool = g.pool()
            ent = self._graphs[key] = (g, new, self._ret)
        ent[0].replay()
        self._new = ent[1]
        self._ret = ent[2]
        return self._ret

    def device_reorder(self) -> None:
        """Queue, right behind ``engine.step``, the gather of every layer's cache rows by the parents the step
        chose (``lgasr_reorder_rows``): afterwards slot ``j`` of utterance ``b`` holds its parent's history."""
        new = self._new
        if not new:  # nothing to re-order (no call yet, or a stateless scorer that re-reads the whole prefix)
            return
        L, Dm = new[0].shape[1], new[0].shape[2]
        if self._buf is None:
            nl, BW = len(new), new[0].shape[0]
            self._buf = [torch.zeros(nl, BW, self._cap, Dm, dtype=new[0].dtype, device=new[0].device) for _ in range(2)]
        dst = self._buf[self._pp ^ 1]
        if isinstance(new, _InPlaceCaches):  # the caches are views of the current buffer: ONE gather for all layers
            self._eng.reorder_rows(dst, self._buf[self._pp][:, :, :L], row_elems=L * Dm)
        else:
            for i, src in enumerate(new):
                self._eng.reorder_rows(dst[i], src.contiguous(), row_elems=L * Dm)
        self._pp ^= 1
        self._new = None
        self._fresh = False


class DeviceCachedDecoderScorer(_DeviceCachedScorer):
    """Full scorer around an attention decoder with the reference's ``BaseTransformerDecoder`` surface
    (``embed``, ``decoders``, ``after_norm``, ``normalize_before``, ``output_layer``; the layers follow the
    ``layer(tgt, tgt_mask, memory, memory_mask, cache=c)`` contract of ``forward_one_step``,
    ``espnet2/asr/decoder/transformer_decoder.py:195-203``).

    Used by this package's ``BatchBeamSearch`` through the device protocol (``device_begin`` /
    ``batch_score_padded`` / ``device_reorder``): all ``B*beam`` row slots are scored in one decoder call per
    position and the caches never leave the device.  Through the plain ``batch_score`` / ``select_state``
    protocol (e.g. under the reference's own search) it simply delegates to the wrapped decoder.
    """

    def __init__(self, decoder: torch.nn.Module, fused_log_softmax: bool = True, cuda_graphs: bool = False,
                 fused_normaliser: bool = False, memory_kv: bool = True):
        for name in ("embed", "decoders", "output_layer"):
            if not hasattr(decoder, name):
                raise TypeError(f"decoder has no attribute {name!r}: not a BaseTransformerDecoder-like module")
        super().__init__(decoder, fused_log_softmax, cuda_graphs, fused_normaliser)
        self.memory_kv = memory_kv

    @property
    def decoder(self):
        return self.net

    def _n_layers(self) -> int:
        return len(self.net.decoders)

    @staticmethod
    def _is_reference_layer(layer) -> bool:
        """The reference's ``DecoderLayer`` (``espnet/nets/pytorch_backend/transformer/decoder_layer.py``): self-attention,
        source attention, feed-forward, three layer norms, no ``concat_after`` / sequential attention."""
        att = getattr(layer, "src_attn", None)
        return (all(hasattr(layer, n) for n in ("self_attn", "src_attn", "feed_forward", "norm1", "norm2", "norm3"))
                and all(hasattr(att, n) for n in ("linear_q", "linear_k", "linear_v", "linear_out", "h", "d_k"))
                and not getattr(layer, "concat_after", False) and getattr(layer, "sequential_attn", None) is None)

    def _begin_inputs(self, xs, lens, B, W, T, dev, fresh_shape):
        layers = list(self.net.decoders)
        # memory_kv: the source-attention keys / values of the encoder memory are projected ONCE per utterance and layer
        # (B x T rows) instead of once per hypothesis row, layer and position (B*beam x T rows, 64 times): the reference
        # layer re-projects the memory inside every call (attention.py forward_qkv), which is ~99 % of its arithmetic
        self._fast = self.memory_kv and all(self._is_reference_layer(l) for l in layers)
        self._BW = (B, W)
        ar = torch.arange(T, device=dev)
        mask = ar[None, :] < torch.as_tensor(list(lens), device=dev)[:, None]  # (B, T)
        if self._fast:
            mem = xs[0].new_zeros(B, T, xs[0].shape[-1])
            for b, x in enumerate(xs):
                mem[b, : lens[b]] = x
            if fresh_shape or getattr(self, "_mem_kv", None) is None:
                # static buffers: the CUDA graphs captured for this shape keep reading them in later batches
                self._mem_kv = [(mem.new_empty(B, l.src_attn.h, l.src_attn.d_k, T), mem.new_empty(B, l.src_attn.h, T, l.src_attn.d_k))
                                for l in layers]
                self._mem_pad = torch.zeros(B, 1, 1, T, dtype=torch.bool, device=dev)
            for l, (kT, v) in zip(layers, self._mem_kv):
                att = l.src_attn
                kT.copy_(att.linear_k(mem).view(B, T, att.h, att.d_k).permute(0, 2, 3, 1))  # K stored transposed: (B, H, dk, T)
                v.copy_(att.linear_v(mem).view(B, T, att.h, att.d_k).transpose(1, 2))
            self._mem_pad.copy_(~mask[:, None, None, :])  # True on padding frames
            return
        if fresh_shape or getattr(self, "_memory", None) is None:
            self._memory = xs[0].new_zeros(B * W, T, xs[0].shape[-1])
            self._memory_mask = torch.zeros(B * W, 1, T, dtype=torch.bool, device=dev)
        mem = self._memory.view(B, W, T, -1)
        mem.zero_()
        for b, x in enumerate(xs):
            mem[b, :, : lens[b]] = x
        self._memory_mask.view(B, W, T).copy_(mask[:, None, :].expand(B, W, T))

    def _layer_fast(self, layer, x, cache, kT, v, tgt_mask, slot=None):
        """One reference ``DecoderLayer`` (``decoder_layer.py:73-150``), same arithmetic per row, with two caches the
        reference does not keep: the memory keys / values of the utterance (``kT``, ``v``; shared by its ``beam`` rows)
        and the layer's own self-attention keys / values of the earlier positions (``cache (n, L-1, 2D)``; the reference
        caches the layer OUTPUTS and re-projects every earlier position at every step, ``attention.py`` forward_qkv).
        ``x``: the layer input, all positions ``(n, L, D)`` without a cache, the last position ``(n, 1, D)`` with one.
        Returns ``(output (n, Lq, D), new cache (n, L, 2D))``."""
        B, W = self._BW
        nb = layer.normalize_before
        sa = layer.self_attn
        n, Lq, D = x.shape
        H, dk = sa.h, sa.d_k
        residual = x
        t = layer.norm1(x) if nb else x
        kv_new = torch.cat([sa.linear_k(t), sa.linear_v(t)], dim=-1)  # (n, Lq, 2D)
        if cache is None:
            kv = kv_new
        elif slot is not None:  # append in place: slot (n, L, 2D) is the persistent buffer the cache is a view of
            slot[:, -1:].copy_(kv_new)
            kv = slot
        else:
            kv = torch.cat([cache, kv_new], dim=1)
        L = kv.shape[1]
        q = sa.linear_q(t).view(n, Lq, H, dk).transpose(1, 2)
        k = kv[..., :D].reshape(n, L, H, dk).transpose(1, 2)
        vv = kv[..., D:].reshape(n, L, H, dk).transpose(1, 2)
        scores = torch.matmul(q, k.transpose(-1, -2)) / (dk ** 0.5)  # (n, H, Lq, L)
        if cache is None and Lq > 1:  # all positions at once: causal mask (the last position alone attends everything)
            m = ~tgt_mask.unsqueeze(1)
            scores = scores.masked_fill(m, torch.finfo(scores.dtype).min)
            p = torch.softmax(scores, dim=-1).masked_fill(m, 0.0)
        else:
            p = torch.softmax(scores, dim=-1)
        h = residual + sa.linear_out(torch.matmul(p, vv).transpose(1, 2).reshape(n, Lq, D))
        if not nb:
            h = layer.norm1(h)
        residual = h
        y = layer.norm2(h) if nb else h
        att = layer.src_attn
        q = att.linear_q(y).view(B, W * Lq, att.h, att.d_k).transpose(1, 2)  # (B, H, W*Lq, dk)
        scores = torch.matmul(q, kT) / (att.d_k ** 0.5)  # (B, H, W*Lq, T)
        scores = scores.masked_fill(self._mem_pad, torch.finfo(scores.dtype).min)
        p = torch.softmax(scores, dim=-1).masked_fill(self._mem_pad, 0.0)
        ctx = torch.matmul(p, v).transpose(1, 2).reshape(n, Lq, D)  # (B, W*Lq, H, dk) -> rows
        h = residual + att.linear_out(ctx)
        if not nb:
            h = layer.norm2(h)
        residual = h
        y = layer.norm3(h) if nb else h
        h = residual + layer.feed_forward(y)
        if not nb:
            h = layer.norm3(h)
        return h, kv

    def _net(self, tgt, cache):
        # the body of forward_one_step (transformer_decoder.py:195-212) without its log_softmax
        dec = self.net
        L = tgt.shape[1]
        tgt_mask = torch.ones(L, L, dtype=torch.bool, device=tgt.device).tril().unsqueeze(0)
        x = dec.embed(tgt)
        new = []
        if self._fast:
            if cache[0] is not None:
                x = x[:, -1:]  # earlier positions live in the key / value caches
            slots = self._slots if (self._slots is not None and cache[0] is not None) else [None] * len(cache)
            for c, layer, (kT, v), sl in zip(cache, dec.decoders, self._mem_kv, slots):
                x, kv = self._layer_fast(layer, x, c, kT, v, tgt_mask, sl)
                new.append(kv)
            self._inplace = slots[0] is not None
        else:
            memory, memory_mask = self._memory, self._memory_mask
            for c, layer in zip(cache, dec.decoders):
                x, tgt_mask, memory, memory_mask = layer(x, tgt_mask, memory, memory_mask, cache=c)
                new.append(x)
        y = x[:, -1]
        if getattr(dec, "normalize_before", False) and getattr(dec, "after_norm", None) is not None:
            y = dec.after_norm(y)
        return dec.output_layer(y), new


class DeviceCachedLMScorer(_DeviceCachedScorer):
    """Full scorer around a language model with the reference's ``TransformerLM`` surface
    (``espnet2/lm/transformer_lm.py:12-133``: ``embed``, ``encoder`` with ``embed`` / ``encoders`` /
    ``normalize_before`` / ``after_norm``, ``decoder``; layers follow ``layer(x, mask, cache=c) -> (x, mask)`` of
    ``Encoder.forward_one_step``, ``espnet/nets/pytorch_backend/transformer/encoder.py:353-379``).  Replaces the
    per-hypothesis state transposes of ``TransformerLM.batch_score`` (``:118-122,132``) the same way."""

    def __init__(self, lm: torch.nn.Module, fused_log_softmax: bool = True, cuda_graphs: bool = False,
                 fused_normaliser: bool = False):
        for name in ("embed", "encoder", "decoder"):
            if not hasattr(lm, name):
                raise TypeError(f"language model has no attribute {name!r}: not a TransformerLM-like module")
        if not hasattr(lm.encoder, "encoders"):
            raise TypeError("language model encoder has no attribute 'encoders'")
        super().__init__(lm, fused_log_softmax, cuda_graphs, fused_normaliser)

    @property
    def lm(self):
        return self.net

    def _n_layers(self) -> int:
        return len(self.net.encoder.encoders)

    def _net(self, tgt, cache):
        lm, enc = self.net, self.net.encoder
        L = tgt.shape[1]
        # _target_mask (:53-56): padding id 0 masked, causal
        mask = (tgt != 0).unsqueeze(-2) & torch.ones(L, L, dtype=torch.bool, device=tgt.device).tril().unsqueeze(0)
        xs = enc.embed(lm.embed(tgt))
        new = []
        for c, layer in zip(cache, enc.encoders):
            xs, mask = layer(xs, mask, cache=c)
            new.append(xs)
        if getattr(enc, "normalize_before", False):
            xs = enc.after_norm(xs)
        return lm.decoder(xs[:, -1]), new


class DevicePrefixLMScorer(_DeviceCachedScorer):
    """Full scorer around a Hugging Face causal language model that re-reads the whole prefix at every position, the
    way the reference's ``HuggingfaceOPTModel.batch_score`` does (``espnet2/lm/huggingface_pretrained_opt_lm.py:110-148``:
    ``self.decoder(ys)`` without a cache, ``lm_head`` on the last position, ``log_softmax``).  Surface: ``lm.decoder``
    (a ``transformers`` base model whose output has ``last_hidden_state``) and ``lm.lm_head``; ``hf_lm.HFPrefixLM`` wraps
    OPT, GPT-2, Llama ... base models that way (GPT-2-small random-init is configs[2]'s shallow-fusion scorer).

    Device protocol: ONE model call per position for all ``B*beam`` row slots (the reference: one per utterance),
    nothing to re-order (stateless), logits handed over through ``lgasr_log_softmax_rows`` or, with
    ``fused_normaliser``, as ``(logits, logsumexp)`` read once by the step in their own type."""

    def __init__(self, lm: torch.nn.Module, fused_log_softmax: bool = True, cuda_graphs: bool = False,
                 fused_normaliser: bool = False):
        for name in ("decoder", "lm_head"):
            if not hasattr(lm, name):
                raise TypeError(f"language model has no attribute {name!r}: not a HuggingfaceOPTModel-like module")
        super().__init__(lm, fused_log_softmax, cuda_graphs, fused_normaliser)

    def _n_layers(self) -> int:
        return 0

    def _net(self, tgt, cache):
        out = self.net.decoder(tgt, use_cache=False, return_dict=True)
        return self.net.lm_head(out.last_hidden_state[:, -1]), []


class DeviceLLMGuidedDecoderScorer(_DeviceCachedScorer):
    """Full scorer around the fork's ``LLMGuidedTransformerDecoder`` (``espnet2/asr/decoder/transformer_decoder.py:944-1237``),
    its ``batch_score`` semantics (``:1115-1132`` -> ``forward_one_step`` ``:1064-1113``): the CTC greedy hypothesis of
    the utterance prompts the LLM (``llm.forward_inference``), the LLM's hidden states of the response tokens go through
    ``embed`` and the attention decoder layers over the encoder memory, ``output_layer`` gives logits over the LLM
    vocabulary (32 k / 128 k entries).

    Surface used: ``dec.ctc.argmax``, ``dec.llm.forward_inference(hyps, hyps_lengths, tgt, tgt_lengths) -> (out, lens)``,
    ``dec.llm.start_of_response_token_id``, ``dec.embed``, ``dec.decoders`` (called on whole sequences, as the reference
    does at ``:1098-1100``), ``dec.after_norm`` / ``normalize_before``, ``dec.output_layer``; optional
    ``dec.ctc_tokenizer`` / ``ctc_token_id_converter`` (text prompts).

    Device protocol: one call per position for all ``B*beam`` row slots of ALL utterances (every row carries its own
    utterance's prompt; the reference handles one utterance at a time and keeps the prompt in ``self.hyp``), logits
    through ``lgasr_log_softmax_rows`` or as ``(logits, logsumexp)`` with ``fused_normaliser`` - the hand-off SURVEY.md
    7.3-8 asks for: the wide bf16 row is read once."""

    def __init__(self, decoder: torch.nn.Module, fused_log_softmax: bool = True, cuda_graphs: bool = False,
                 fused_normaliser: bool = False):
        for name in ("embed", "decoders", "output_layer", "llm", "ctc"):
            if getattr(decoder, name, None) is None:
                raise TypeError(f"decoder has no {name!r}: not a configured LLMGuidedTransformerDecoder")
        super().__init__(decoder, fused_log_softmax, cuda_graphs, fused_normaliser)

    def _n_layers(self) -> int:
        return 0  # batch_score (use_cache False) keeps no state: the whole prefix is re-read

    def _begin_inputs(self, xs, lens, B, W, T, dev, fresh_shape):
        dec = self.net
        if fresh_shape:
            self._memory = xs[0].new_zeros(B * W, T, xs[0].shape[-1])
        mem = self._memory.view(B, W, T, -1)
        mem.zero_()
        hyps = []
        for b, x in enumerate(xs):
            mem[b, :, : lens[b]] = x
            lpz = dec.ctc.argmax(x.unsqueeze(0)).data  # :1071-1074: greedy CTC, collapse repeats, drop blanks
            y_hat = torch.unique_consecutive(lpz[0])
            hyp = y_hat[y_hat != 0]
            if getattr(dec, "ctc_tokenizer", None) is not None:
                hyp = dec.ctc_tokenizer.tokens2text(dec.ctc_token_id_converter.ids2tokens(hyp))
            hyps.append(hyp)
        self._hyps = [hyps[b] for b in range(B) for _ in range(W)]
        text = getattr(dec, "ctc_tokenizer", None) is not None
        self._hyp_lens = torch.tensor([-1 if text else int(h.shape[0]) for h in self._hyps], dtype=torch.int64, device=dev)
        # the reference passes memory_mask None (:1099): every frame of the (single) utterance is attended.  In a batch
        # the shorter utterances are padded, so their padding frames are masked out
        ar = torch.arange(T, device=dev)
        mask = ar[None, :] < torch.as_tensor(list(lens), device=dev)[:, None]
        self._memory_mask = mask[:, None, None, :].expand(B, W, 1, T).reshape(B * W, 1, T) if min(lens) < T else None

    def _net(self, tgt, cache):
        dec = self.net
        n, L = tgt.shape
        tgt = tgt.clone()
        tgt[:, 0] = dec.llm.start_of_response_token_id  # :1087
        llm_out, _ = dec.llm.forward_inference(self._hyps, self._hyp_lens, tgt, tgt.new_full((n,), L))
        tgt_mask = torch.ones(L, L, dtype=torch.bool, device=tgt.device).tril().unsqueeze(0)
        x = dec.embed(llm_out)
        x, _, _, _ = dec.decoders(x, tgt_mask, self._memory, self._memory_mask)
        y = dec.after_norm(x[:, -1]) if getattr(dec, "normalize_before", False) else x[:, -1]
        return dec.output_layer(y), []
