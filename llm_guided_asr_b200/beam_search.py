"""Drop-in ``BatchBeamSearch`` whose per-position scoring runs as hand-written sm_100a CUDA.

API mirror of the reference search objects:
``Hypothesis`` (``espnet/nets/beam_search.py:15-31``), ``BeamSearch.__init__`` (``:37-121``),
``BeamSearch.forward`` (``:385-498``) and ``BatchBeamSearch`` (``espnet/nets/batch_beam_search.py``):
same constructor arguments, same call ``beam_search(x=enc (T, D), maxlenratio, minlenratio)``,
same result (n-best list of ``Hypothesis``, best first), same errors for bad scorers / keys.

What differs is *where* the work happens.  The reference runs, per decode position, ~2 k tiny
eager ops plus O(beam) device->host syncs.  Here the full scorers
(attention decoder, LM, LLM - producers of ``(n, V)`` log-prob matrices) still run
in PyTorch, and everything after them - weighted fusion, pre-beam top-k, the CTC prefix forward
recursion, global top-k, hypothesis bookkeeping, eos handling and end detection - is two CUDA
kernels behind the C ABI (``include/lgasr.h``), with one small D2H read per position.

There is no CPU path: construction works anywhere, decoding needs the CUDA library and a GPU.
"""

from collections import OrderedDict
from typing import Any, Dict, List, NamedTuple, Optional, Sequence, Union

import torch

from .engine import FusedStepEngine
from .scorer_interface import BatchScorerInterface, PartialScorerInterface, ScorerInterface


class Hypothesis(NamedTuple):
    """Result type, field-compatible with ``espnet.nets.beam_search.Hypothesis``."""

    yseq: torch.Tensor
    score: Union[float, torch.Tensor] = 0
    scores: Dict[str, Union[float, torch.Tensor]] = dict()
    states: Dict[str, Any] = dict()
    hs: List[torch.Tensor] = []

    def asdict(self) -> dict:
        return self._replace(
            yseq=self.yseq.tolist(),
            score=float(self.score),
            scores={k: float(v) for k, v in self.scores.items()},
        )._asdict()


def length_bounds(T: int, maxlenratio: float, minlenratio: float):
    """maxlen / minlen of ``BeamSearch.forward`` (``beam_search.py:417-427``)."""
    if maxlenratio == 0:
        maxlen = T
    elif maxlenratio < 0:
        maxlen = -1 * int(maxlenratio)
    else:
        maxlen = max(1, int(maxlenratio * T))
    minlen = -1 * int(minlenratio) if minlenratio < 0 else int(minlenratio * T)
    return maxlen, minlen


def _auto_device_cache(v):
    """Device-protocol wrapper for a full scorer with the reference's ``BaseTransformerDecoder`` or ``TransformerLM``
    surface, None for anything else (scorers that already speak the device protocol, table scorers, the LLM-guided
    decoder and HF language models, which the caller wraps explicitly: their prompts / tokenizers need arguments)."""
    if hasattr(v, "batch_score_padded") or not isinstance(v, torch.nn.Module) or hasattr(v, "llm"):
        return None
    from .device_cache import DeviceCachedDecoderScorer, DeviceCachedLMScorer

    if all(hasattr(v, a) for a in ("embed", "decoders", "output_layer", "forward_one_step")):
        return DeviceCachedDecoderScorer(v)
    if all(hasattr(v, a) for a in ("embed", "encoder", "decoder")) and hasattr(v.encoder, "encoders"):
        return DeviceCachedLMScorer(v)
    return None


class BatchBeamSearch(torch.nn.Module):
    """Fused batch beam search (one or many utterances in lock-step)."""

    def __init__(
        self,
        scorers: Dict[str, ScorerInterface],
        weights: Dict[str, float],
        beam_size: int,
        vocab_size: int,
        sos: int,
        eos: int,
        token_list: List[str] = None,
        pre_beam_ratio: float = 1.5,
        pre_beam_score_key: str = None,
        return_hs: bool = False,
        hyp_primer: List[int] = None,
        normalize_length: bool = False,
        device_caches: bool = True,
    ):
        super().__init__()
        self.weights = weights
        self.scorers = dict()
        self.full_scorers = dict()
        self.part_scorers = dict()
        self.nn_dict = torch.nn.ModuleDict()
        for k, v in scorers.items():
            w = weights.get(k, 0)
            if w == 0 or v is None:  # dropped entirely, beam_search.py:82-85
                continue
            assert isinstance(v, ScorerInterface), f"{k} ({type(v)}) does not implement ScorerInterface"
            self.scorers[k] = v
            if isinstance(v, PartialScorerInterface):
                self.part_scorers[k] = v
            else:
                self.full_scorers[k] = v
            if isinstance(v, torch.nn.Module):
                self.nn_dict[k] = v
        for k, v in self.full_scorers.items():
            if not isinstance(v, BatchScorerInterface):
                raise NotImplementedError(f"full scorer {k} is not a BatchScorerInterface (asr_inference.py:365-380)")
        if device_caches:
            # the reference's own attention decoder / TransformerLM handed over as they are (the plain class swap of
            # INTEGRATION.md): wrap them so that decode_batch calls them once per position for all row slots and keeps
            # their caches on the device, instead of one batch_score per utterance with per-hypothesis state lists
            # (batch_beam_search.py:347-355, transformer_decoder.py:268-272,285-286).  device_caches=False keeps
            # the generic protocol.
            for k, v in list(self.full_scorers.items()):
                w = _auto_device_cache(v)
                if w is not None:
                    self.full_scorers[k] = self.scorers[k] = w
                    self.nn_dict[k] = w
        if len(self.part_scorers) > 1 or any(not hasattr(v, "ctc") for v in self.part_scorers.values()):
            raise NotImplementedError("the fused step supports exactly one partial scorer: the CTC prefix scorer")
        if return_hs:
            raise NotImplementedError("return_hs is not supported by the fused search")
        self.sos, self.eos = sos, eos
        self.hyp_primer = hyp_primer
        self.token_list = token_list
        self.pre_beam_size = int(pre_beam_ratio * beam_size)
        self.beam_size = beam_size
        self.n_vocab = vocab_size
        if (
            pre_beam_score_key is not None
            and pre_beam_score_key != "full"
            and pre_beam_score_key not in self.full_scorers
        ):
            raise KeyError(f"{pre_beam_score_key} is not found in {self.full_scorers}")
        self.pre_beam_score_key = pre_beam_score_key
        self.do_pre_beam = (
            self.pre_beam_score_key is not None
            and self.pre_beam_size < self.n_vocab
            and len(self.part_scorers) > 0
        )
        self.return_hs = return_hs
        self.normalize_length = normalize_length
        self._engines: "OrderedDict[tuple, FusedStepEngine]" = OrderedDict()
        self.last_steps = 0  # decode positions executed by the last call
        # poll_every: positions between two host reads of the done flags when every full scorer keeps its state on the
        # device (nothing on the host depends on the step's outcome then; positions run past an utterance's end cost
        # idle CTAs only).  None: 1 with end detection (maxlenratio == 0), 8 otherwise.
        self.poll_every: Optional[int] = None
        self.collect_timing = False  # True: CUDA-event split of decode_batch into scorer calls / step / reorder
        self.last_timing: Dict[str, float] = {}

    def set_hyp_primer(self, hyp_primer: List[int] = None) -> None:
        self.hyp_primer = hyp_primer

    # ------------------------------------------------------------------------------------------
    def _engine(self, B, T, maxcap, primer_len, end_detect, device) -> FusedStepEngine:
        """Device workspace for a batch: cached on BUCKETED capacities (frames and positions rounded up to 64), so
        that utterances of similar length reuse one handle (per-utterance ``xlens`` / ``maxlen`` go to
        ``set_posteriors`` / ``begin``); least-recently-used eviction."""
        Tcap = -(-T // 64) * 64
        Lcap = -(-maxcap // 64) * 64
        key = (B, Tcap, Lcap, primer_len, end_detect, str(device))
        eng = self._engines.get(key)
        if eng is not None:
            self._engines.move_to_end(key)
        else:
            if len(self._engines) >= 4:  # bounded cache of device workspaces: drop the least recently used
                _, old = self._engines.popitem(last=False)
                old.close()
            T, maxcap = Tcap, Lcap
            ctc_name = next(iter(self.part_scorers), None)
            eng = FusedStepEngine(
                n_utts=B,
                max_frames=T,
                vocab=self.n_vocab,
                beam=self.beam_size,
                pre_beam=self.pre_beam_size,
                term_names=list(self.full_scorers),
                pre_beam_key=self.pre_beam_score_key,
                use_ctc=ctc_name is not None,
                ctc_weight=self.weights[ctc_name] if ctc_name else 0.0,
                sos=self.sos,
                eos=self.eos,
                max_len_cap=maxcap,
                primer_len=primer_len,
                end_detect=end_detect,
                device=device,
            )
            self._engines[key] = eng
        return eng

    def forward(
        self, x: torch.Tensor, maxlenratio: float = 0.0, minlenratio: float = 0.0, pre_x: torch.Tensor = None
    ) -> List[Hypothesis]:
        """One utterance: ``x (T, D)`` encoder output -> n-best ``Hypothesis`` list, best first."""
        if pre_x is not None:
            raise NotImplementedError("sequential attention (pre_x) is outside the fused path")
        return self.decode_batch([x], maxlenratio, minlenratio)[0]

    def _hyps_of(self, eng, b, out_names, device) -> List[Hypothesis]:
        """N-best list of slot ``b`` (ended hypotheses, best first; ``beam_search.py:452-459``)."""
        toks, score, scores = eng.gather_ended(b)
        hyps = [
            Hypothesis(
                yseq=torch.tensor(t, dtype=torch.int64, device=device),
                score=torch.tensor(float(score[e]), dtype=torch.float32, device=device),
                scores={k: torch.tensor(float(scores[e, j]), device=device) for j, k in enumerate(out_names)},
                states={},
            )
            for e, t in enumerate(toks)
        ]
        keys = [float(score[e]) / (len(t) - 1) if self.normalize_length else float(score[e]) for e, t in enumerate(toks)]
        return [hyps[i] for i in sorted(range(len(hyps)), key=lambda i: keys[i], reverse=True)]

    @torch.no_grad()
    def decode_stream(
        self, xs: Sequence[torch.Tensor], slots: int, maxlenratio: float = 0.0, minlenratio: float = 0.0
    ) -> List[List[Hypothesis]]:
        """Continuous batching: ``len(xs)`` utterances through ``slots`` lock-step slots; the moment an utterance ends
        (``done`` flag of the step: end detection, no running hypothesis left, or its maximum length) its slot is
        handed to the next utterance of the list (``lgasr_refill``) while the others keep decoding, so no slot idles
        behind the longest utterance of a batch.  Per utterance identical to ``forward``.

        Full scorers are driven through the plain ``BatchScorerInterface`` protocol, one call per utterance and
        position (the slots are at different prefix lengths)."""
        n_utt = len(xs)
        if n_utt == 0:
            return []
        B = min(int(slots), n_utt)
        device = xs[0].device
        if device.type != "cuda":
            raise RuntimeError("BatchBeamSearch (fused, sm_100a) needs CUDA tensors: there is no CPU fallback")
        lens = [int(x.shape[0]) for x in xs]
        bounds = [length_bounds(t, maxlenratio, minlenratio) for t in lens]
        primer = [self.sos] if self.hyp_primer is None else list(self.hyp_primer)
        P = len(primer)
        eng = self._engine(B, max(lens), max(b[0] for b in bounds), P, maxlenratio == 0.0, device)
        W, V = self.beam_size, self.n_vocab
        ctc_name = next(iter(self.part_scorers), None)
        ctc = self.part_scorers[ctc_name].ctc if ctc_name is not None else None

        def posteriors(u):
            lp = torch.empty(eng.T, V, dtype=torch.float32, device=device)
            lp[: lens[u]] = ctc.log_softmax(xs[u].unsqueeze(0))[0].to(torch.float32)
            return lp

        slot_utt = list(range(B))  # utterance held by each slot
        nxt = B
        if ctc is not None:
            eng.set_posteriors(torch.stack([posteriors(u) for u in slot_utt]), [lens[u] for u in slot_utt])
        else:
            eng.set_posteriors(None, [lens[u] for u in slot_utt])
        eng.begin([bounds[u][0] for u in slot_utt], [bounds[u][1] for u in slot_utt], primer)
        names = list(self.full_scorers)
        wts = [float(self.weights[k]) for k in names]
        out_names = names + ([ctc_name] if ctc_name is not None else [])
        states = {k: [[self.full_scorers[k].batch_init_state(xs[u])] for u in slot_utt] for k in names}
        n_run, done, cur_len = [1] * B, [0] * B, [P] * B
        pad = {k: None for k in names}
        results: List[Optional[List[Hypothesis]]] = [None] * n_utt
        steps = 0
        while True:
            yseq = eng.view(0)
            mats, consts = [], []
            new_states = {k: [None] * B for k in names}
            for k in names:
                d = self.full_scorers[k]
                if getattr(d, "constant_score", None) is not None:
                    mats.append(None)
                    consts.append(float(d.constant_score))
                    continue
                if pad[k] is None:
                    pad[k] = torch.zeros(B * W, V, dtype=torch.float32, device=device)
                for b in range(B):
                    if done[b] or n_run[b] == 0:
                        continue
                    u = slot_utt[b]
                    ys = yseq[b, : n_run[b], : cur_len[b]]
                    sc, st = d.batch_score(ys, states[k][b], xs[u].expand(n_run[b], *xs[u].shape))
                    new_states[k][b] = st
                    pad[k][b * W : b * W + n_run[b]] = sc
                mats.append(pad[k])
                consts.append(0.0)
            eng.step(mats, consts, wts)
            steps += 1
            nrun_new, done_new, par, _tok, _act = eng.poll(with_parents=True)
            for b in range(B):
                if done[b]:
                    continue
                cur_len[b] += 1
                for k in names:
                    d = self.full_scorers[k]
                    if new_states[k][b] is not None:
                        states[k][b] = [d.select_state(new_states[k][b], int(par[b, r])) for r in range(int(nrun_new[b]))]
                n_run[b] = int(nrun_new[b])
                if int(done_new[b]):
                    u = slot_utt[b]
                    hyps = self._hyps_of(eng, b, out_names, device)
                    if len(hyps) == 0 and minlenratio >= 0.1:  # beam_search.py:462-471
                        hyps = self.decode_batch([xs[u]], maxlenratio, max(0.0, minlenratio - 0.1))[0]
                    results[u] = hyps
                    if nxt < n_utt:  # the slot takes the next utterance at once
                        u2 = nxt
                        nxt += 1
                        eng.refill(b, posteriors(u2) if ctc is not None else None, lens[u2], bounds[u2][0], bounds[u2][1])
                        slot_utt[b] = u2
                        for k in names:
                            states[k][b] = [self.full_scorers[k].batch_init_state(xs[u2])]
                        n_run[b], cur_len[b] = 1, P
                    else:
                        done[b] = 1
            if all(done):
                break
        self.last_steps = steps
        return results

    @torch.no_grad()
    def decode_batch(
        self, xs: Sequence[torch.Tensor], maxlenratio: float = 0.0, minlenratio: float = 0.0
    ) -> List[List[Hypothesis]]:
        """Many utterances in lock-step: ``xs[b] (T_b, D)``.  Per utterance identical to ``forward``."""
        B = len(xs)
        device = xs[0].device
        if device.type != "cuda":
            raise RuntimeError("BatchBeamSearch (fused, sm_100a) needs CUDA tensors: there is no CPU fallback")
        lens = [int(x.shape[0]) for x in xs]
        T = max(lens)
        bounds = [length_bounds(t, maxlenratio, minlenratio) for t in lens]
        maxlen = [b[0] for b in bounds]
        minlen = [b[1] for b in bounds]
        primer = [self.sos] if self.hyp_primer is None else list(self.hyp_primer)
        eng = self._engine(B, T, max(maxlen), len(primer), maxlenratio == 0.0, device)
        W, V = self.beam_size, self.n_vocab

        # CTCPrefixScorer.batch_init_state: logp = ctc.log_softmax(x[None]) (scorers/ctc.py:96)
        ctc_name = next(iter(self.part_scorers), None)
        if ctc_name is not None:
            ctc = self.part_scorers[ctc_name].ctc
            # frames past an utterance's length are never read (lgasr_set_posteriors applies the padding rule of
            # ctc_prefix_score.py:46-49 itself), so the batch buffer is not cleared; one CTC-head call per utterance
            # keeps every utterance's posteriors bit-identical to a single-utterance decode
            logp = torch.empty(B, eng.T, V, dtype=torch.float32, device=device)
            for b, x in enumerate(xs):
                logp[b, : lens[b]] = ctc.log_softmax(x.unsqueeze(0))[0].to(torch.float32)
            eng.set_posteriors(logp, lens)
            del logp
        else:
            eng.set_posteriors(None, lens)
        eng.begin(maxlen, minlen, primer)

        names = list(self.full_scorers)
        wts = [float(self.weights[k]) for k in names]
        # per-utterance scorer state lists (init_hyp, batch_beam_search.py:142-144)
        padded = {k for k in names if hasattr(self.full_scorers[k], "batch_score_padded")}
        states = {k: [[self.full_scorers[k].batch_init_state(x)] for x in xs] for k in names if k not in padded}
        for k in names:  # scorers that keep their per-hypothesis state on the device (device_cache.py)
            if hasattr(self.full_scorers[k], "device_begin"):
                self.full_scorers[k].device_begin(xs, lens, eng, len(primer) + max(maxlen))
        n_run = [1] * B
        done = [0] * B
        pad = {k: None for k in names}
        steps = 0
        all_device = len(padded) + sum(1 for k in names if getattr(self.full_scorers[k], "constant_score", None) is not None) == len(names)
        every = self.poll_every if self.poll_every is not None else (1 if maxlenratio == 0.0 else 8)
        every = max(1, every) if all_device else 1
        ev = [] if self.collect_timing else None

        def mark(tag):
            if ev is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                ev.append((tag, e))

        for i in range(max(maxlen)):
            cur_len = len(primer) + i
            yseq = eng.view(0)  # (B, W, Lcap) current running prefixes
            mark("start")
            mats, consts, norm_at = [], [], {}
            new_states = {k: [None] * B for k in names}
            for k in names:
                d = self.full_scorers[k]
                if getattr(d, "constant_score", None) is not None:  # LengthBonus: folded as a scalar term
                    mats.append(None)
                    consts.append(float(d.constant_score))
                    continue
                if hasattr(d, "batch_score_padded"):  # scorer that fills the padded row layout itself
                    res = d.batch_score_padded(yseq[:, :, :cur_len], n_run, done)
                    if isinstance(res, tuple):  # (logits, per-row logsumexp): normalised inside the step kernel
                        mats.append(res[0])
                        norm_at[len(mats) - 1] = res[1]
                    else:
                        mats.append(res)
                    consts.append(0.0)
                    continue
                if B == 1 and not done[0]:
                    ys = yseq[0, : n_run[0], :cur_len]
                    sc, st = d.batch_score(ys, states[k][0], xs[0].expand(n_run[0], *xs[0].shape))
                    new_states[k][0] = st
                    sc = sc.to(torch.float32)
                    if sc.stride(-1) != 1 or (sc.shape[0] > 1 and sc.stride(0) < V):
                        if sc.stride(0) == 0 and sc.stride(-1) == 0:  # constant row (LengthBonus expand)
                            mats.append(None)
                            consts.append(float(sc.flatten()[0]))
                            continue
                        sc = sc.contiguous()
                    mats.append(sc)
                    consts.append(0.0)
                    continue
                # generic multi-utterance path: score utterance by utterance, scatter into padded rows
                if pad[k] is None:
                    pad[k] = torch.zeros(B * W, V, dtype=torch.float32, device=device)
                const_val = None
                for b in range(B):
                    if done[b] or n_run[b] == 0:
                        continue
                    ys = yseq[b, : n_run[b], :cur_len]
                    sc, st = d.batch_score(ys, states[k][b], xs[b].expand(n_run[b], *xs[b].shape))
                    new_states[k][b] = st
                    if sc.stride(0) == 0 and sc.stride(-1) == 0:
                        const_val = float(sc.flatten()[0])
                    else:
                        pad[k][b * W : b * W + n_run[b]] = sc
                if const_val is not None:
                    mats.append(None)
                    consts.append(const_val)
                else:
                    mats.append(pad[k])
                    consts.append(0.0)
            mark("scorers")
            eng.step(mats, consts, wts, [norm_at.get(j) for j in range(len(mats))] if norm_at else None)
            mark("step")
            steps += 1
            for k in names:  # device-resident caches: parent-indexed gather queued behind the step
                if hasattr(self.full_scorers[k], "device_reorder"):
                    self.full_scorers[k].device_reorder()
            mark("reorder")
            if all_device and (i + 1) % every != 0 and i + 1 < max(maxlen):
                continue  # no host read this position: the next scorer call is queued right behind the gather
            nrun_new, done_new, par, _tok, n_active = eng.poll(with_parents=len(padded) < len(names))
            # re-order the full scorers' states by the parent slots (batch_beam_search.py:347-351)
            for k in names:
                if k in padded:
                    continue
                d = self.full_scorers[k]
                for b in range(B):
                    if done[b]:
                        continue
                    st = new_states[k][b]
                    states[k][b] = [d.select_state(st, int(par[b, r])) for r in range(int(nrun_new[b]))]
            n_run = [int(v) for v in nrun_new]
            done = [int(v) for v in done_new]
            if n_active == 0:
                break
        self.last_steps = steps
        if ev is not None:
            torch.cuda.synchronize(device)
            acc = {"scorers": 0.0, "step": 0.0, "reorder": 0.0}
            for (t0, e0), (t1, e1) in zip(ev, ev[1:]):
                if t1 in acc:
                    acc[t1] += e0.elapsed_time(e1)
            self.last_timing = dict(acc, positions=steps)

        results = []
        out_names = names + ([ctc_name] if ctc_name is not None else [])
        # every hypothesis of the batch in THREE host->device copies (tokens, scores, per-scorer scores); the Hypothesis
        # fields are views of those tensors
        gathered = [eng.gather_ended(b) for b in range(B)]
        flat_tok = [t for toks, _, _ in gathered for seq in toks for t in seq]
        tok_dev = torch.tensor(flat_tok, dtype=torch.int64, device=device) if flat_tok else torch.zeros(0, dtype=torch.int64, device=device)
        n_h = sum(len(toks) for toks, _, _ in gathered)
        NSo = max(len(out_names), 1)
        sc_host = torch.zeros(n_h, 1 + NSo, dtype=torch.float32)
        r = 0
        for toks, score, scores in gathered:
            for e in range(len(toks)):
                sc_host[r, 0] = float(score[e])
                for j in range(len(out_names)):
                    sc_host[r, 1 + j] = float(scores[e, j])
                r += 1
        sc_dev = sc_host.to(device)
        off, r = 0, 0
        for b in range(B):
            toks, score, scores = gathered[b]
            hyps, keys = [], []
            for e, t in enumerate(toks):
                hyps.append(Hypothesis(
                    yseq=tok_dev[off : off + len(t)],
                    score=sc_dev[r, 0],
                    scores={k: sc_dev[r, 1 + j] for j, k in enumerate(out_names)},
                    states={},
                ))
                # sort keys from the host copy (no device read per hypothesis): python floats of the fp32 score,
                # / (len(yseq) - 1) when normalising (beam_search.py:452-459)
                sc = float(score[e])
                keys.append(sc / (len(t) - 1) if self.normalize_length else sc)
                off += len(t)
                r += 1
            order = sorted(range(len(hyps)), key=lambda i: keys[i], reverse=True)  # python's sort is stable
            hyps = [hyps[i] for i in order]
            if len(hyps) == 0:  # retry with a smaller minlenratio, beam_search.py:462-471
                if minlenratio >= 0.1:
                    hyps = self.decode_batch([xs[b]], maxlenratio, max(0.0, minlenratio - 0.1))[0]
            results.append(hyps)
        return results
