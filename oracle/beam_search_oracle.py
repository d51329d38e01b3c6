"""ORACLE (test infrastructure, not product code): CPU restatement of the batched beam search.

NumPy fp32 restatement of the reference's joint CTC/attention(/LM) beam search for one
utterance: ``BeamSearch.__init__`` scorer partition (``espnet/nets/beam_search.py:37-121``),
``BeamSearch.forward`` (``:385-498``), ``BatchBeamSearch.search`` / ``batch_beam`` / ``post_process``
(``espnet/nets/batch_beam_search.py:101-128,259-363,365-429``) and ``end_detect``
(``espnet/nets/e2e_asr_common.py:17-48``), driving ``oracle.ctc_prefix_oracle.CTCPrefixOracle``.
Utterances are decoded one at a time, as the reference requires
(``espnet2/bin/asr_inference.py:759-764``).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import this.
Pinned against the reference's own n-best lists in ``tests/golden/*.npz``
(``tests/test_oracle_golden.py``).

Top-k tie rule (the reference's ``torch.topk`` order among equal values is unspecified,
SURVEY.md §3.4-6): higher score first, then lower flat index — the same rule the CUDA path uses.
"""

from typing import Callable, Dict, List, Optional

import numpy as np

from .ctc_prefix_oracle import CTCPrefixOracle

D_END = float(np.log(1 * np.exp(-10)))  # e2e_asr_common.py:17


def topk_desc(values: np.ndarray, k: int) -> np.ndarray:
    """Indices of the k largest entries of a 1-D array, descending, ties by lower index."""
    order = np.argsort(-values.astype(np.float64), kind="stable")
    return order[:k]


def end_detect(ended, i, M=3, d_end=D_END):
    """``ended``: list of (score float, len(yseq) int).  e2e_asr_common.py:17-48."""
    if len(ended) == 0:
        return False
    best = max(s for s, _ in ended)
    count = 0
    for m in range(M):
        same = [s for s, ln in ended if ln == i - m]
        if len(same) > 0 and max(same) - best < d_end:
            count += 1
    return count == M


class Hyp:
    __slots__ = ("yseq", "score", "scores", "ctc_r", "ctc_s")

    def __init__(self, yseq, score, scores, ctc_r=None, ctc_s=None):
        self.yseq, self.score, self.scores = yseq, score, scores
        self.ctc_r, self.ctc_s = ctc_r, ctc_s


class BeamSearchOracle:
    """``full_scorers``: ordered dict name -> callable(ys int64 (n, L)) -> (n, V) fp32 array.
    ``weights``: name -> float, including ``"ctc"`` when a CTC scorer is used."""

    def __init__(
        self,
        full_scorers: Dict[str, Callable],
        weights: Dict[str, float],
        beam_size: int,
        vocab_size: int,
        sos: int,
        eos: int,
        blank: int = 0,
        pre_beam_ratio: float = 1.5,
        pre_beam_score_key: Optional[str] = None,
        normalize_length: bool = False,
        hyp_primer: Optional[List[int]] = None,
        use_ctc: bool = True,
    ):
        # scorers with weight 0 are dropped entirely (beam_search.py:82-85)
        self.full = {k: f for k, f in full_scorers.items() if weights.get(k, 0) != 0}
        self.use_ctc = use_ctc and weights.get("ctc", 0) != 0
        self.w = {k: np.float32(v) for k, v in weights.items()}
        self.beam, self.V, self.sos, self.eos, self.blank = beam_size, vocab_size, sos, eos, blank
        self.S = int(pre_beam_ratio * beam_size)  # :105
        if pre_beam_score_key is not None and pre_beam_score_key != "full" and pre_beam_score_key not in self.full:
            raise KeyError(pre_beam_score_key)  # :108-113
        self.key = pre_beam_score_key
        self.do_pre_beam = self.key is not None and self.S < self.V and self.use_ctc  # :115-119
        self.normalize_length = normalize_length
        self.primer = hyp_primer
        self.names = list(self.full) + (["ctc"] if self.use_ctc else [])
        self.trace = None  # optional list collecting per-step records

    # -- one decode position: batch_beam_search.py:259-363 ---------------------------------
    def search(self, running: List[Hyp], ctc: Optional[CTCPrefixOracle]) -> List[Hyp]:
        n, V = len(running), self.V
        ys = np.array([h.yseq for h in running], dtype=np.int64)
        weighted = np.zeros((n, V), dtype=np.float32)
        scores = {}
        for k, f in self.full.items():
            scores[k] = np.asarray(f(ys), dtype=np.float32)
            weighted += self.w[k] * scores[k]
        part_ids = None
        if self.do_pre_beam:
            pre = weighted if self.key == "full" else scores[self.key]
            part_ids = np.stack([topk_desc(pre[h], self.S) for h in range(n)])
        part = None
        if self.use_ctc:
            r_prev = np.stack([h.ctc_r for h in running], axis=2)
            s_prev = np.array([h.ctc_s for h in running], dtype=np.float32)
            last = [h.yseq[-1] for h in running]
            part, state = ctc(last, ys.shape[1] - 1, r_prev, s_prev, part_ids)
            weighted += self.w["ctc"] * part
        weighted += np.array([h.score for h in running], dtype=np.float32)[:, None]
        top = topk_desc(weighted.reshape(-1), self.beam)  # :101-128
        best = []
        for flat in top:
            h, tok = int(flat // V), int(flat % V)
            p = running[h]
            sc = {k: np.float32(p.scores[k] + scores[k][h, tok]) for k in self.full}
            r_new = s_new = None
            if self.use_ctc:
                sc["ctc"] = np.float32(p.scores["ctc"] + part[h, tok])
                r_new, s_new = CTCPrefixOracle.select_state(state, h, tok)
            best.append(Hyp(p.yseq + [tok], weighted[h, tok], sc, r_new, s_new))
        if self.trace is not None:
            self.trace.append(
                dict(
                    ys=ys,
                    part_ids=part_ids,
                    ctc_scores=part,
                    top_hyp=np.array([int(f // V) for f in top]),
                    top_tok=np.array([int(f % V) for f in top]),
                    top_score=np.array([weighted.reshape(-1)[f] for f in top], dtype=np.float32),
                )
            )
        return best

    # -- batch_beam_search.py:365-429 ------------------------------------------------------
    def post_process(self, i, maxlen, minlen, best: List[Hyp], ended: List[Hyp]) -> List[Hyp]:
        if i == maxlen - 1:
            for h in best:
                h.yseq = h.yseq + [self.eos]
        remained = []
        for h in best:
            if h.yseq[-1] == self.eos:
                if i >= minlen:
                    ended.append(h)
            else:
                remained.append(h)
        return remained

    # -- beam_search.py:385-498 ------------------------------------------------------------
    def forward(self, logp: Optional[np.ndarray], maxlenratio=0.0, minlenratio=0.0, T: Optional[int] = None):
        """``logp (T, V)`` CTC log-posteriors (None when no CTC scorer; then give ``T``)."""
        T = logp.shape[0] if logp is not None else T
        if maxlenratio == 0:
            maxlen = T
        elif maxlenratio < 0:
            maxlen = -1 * int(maxlenratio)
        else:
            maxlen = max(1, int(maxlenratio * T))
        minlen = -1 * int(minlenratio) if minlenratio < 0 else int(minlenratio * T)

        ctc = CTCPrefixOracle(logp, self.blank, self.eos) if self.use_ctc else None
        primer = [self.sos] if self.primer is None else list(self.primer)
        init = Hyp(primer, np.float32(0.0), {k: np.float32(0.0) for k in self.names})
        if self.use_ctc:
            r0, s0 = ctc.initial_state()
            init.ctc_r, init.ctc_s = r0[:, :, 0], s0[0]
        running, ended = [init], []
        self.steps_run = 0
        for i in range(maxlen):
            best = self.search(running, ctc)
            self.steps_run += 1
            running = self.post_process(i, maxlen, minlen, best, ended)
            if maxlenratio == 0.0 and end_detect([(float(h.score), len(h.yseq)) for h in ended], i):
                break
            if len(running) == 0:
                break
        if self.normalize_length:
            key = lambda h: float(h.score) / (len(h.yseq) - 1)  # noqa: E731
        else:
            key = lambda h: float(h.score)  # noqa: E731
        nbest = sorted(ended, key=key, reverse=True)
        if len(nbest) == 0:  # :462-471
            if minlenratio < 0.1:
                return []
            return self.forward(logp, maxlenratio, max(0.0, minlenratio - 0.1), T=T)
        return nbest


def torch_scorer_fn(scorer):
    """Wrap a torch ``BatchScorerInterface`` object (CPU) as an oracle full-scorer callable."""
    import torch

    def fn(ys):
        out, _ = scorer.batch_score(torch.as_tensor(ys), [None] * len(ys), None)
        return out.detach().cpu().numpy()

    return fn
