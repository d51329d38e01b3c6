"""Drop-in ``CTCPrefixScorer`` (API mirror of ``espnet/nets/scorers/ctc.py:10-157``) on CUDA.

Same constructor and methods as the reference class, so an *unmodified* reference
``BatchBeamSearch`` (or the non-batch ``BeamSearch``) can drive it through the plug-in API:

* ``batch_init_state(x)``: ``logp = ctc.log_softmax(x[None])`` then the posteriors are re-laid out
  frames-contiguous on the device (replaces ``CTCPrefixScoreTH.__init__``, ``ctc_prefix_score.py:21-65``)
* ``batch_score_partial(y, ids, state, x)`` -> ``(scores (n, V), state)``: one kernel computes
  ``log psi`` of the candidates by warp log-sum-exp (replaces ``CTCPrefixScoreTH.__call__`` ``:67-187``)
* ``select_state(state, i, new_id)``: the forward recursion of the survivor only (replaces the
  ``r[:, :, i, idmap[i, new_id]]`` pick, ``scorers/ctc.py:56-62``, of an r tensor the reference
  materialises for every candidate)

The fused search (``llm_guided_asr_b200.beam_search.BatchBeamSearch``) does not call these methods:
it only takes ``.ctc`` and ``.eos`` from this object and runs the same device functions inside the
fused step kernels.  ``extend_prob`` / ``extend_state`` are the streaming extension of the scorer API
(``scorers/ctc.py:128-157``); the per-hypothesis state here is ``(r (2, Tp), s, T, (f_min, f_max))``.
"""

import ctypes as C
from typing import Any, List, Tuple

import torch

from . import _lib
from .scorer_interface import BatchPartialScorerInterface

LOGZERO = -10000000000.0


def _st(dev):
    return int(torch.cuda.current_stream(dev).cuda_stream)


def _index_vector(vals, dev):
    """int64 device vector from Python ints or 0-dim tensors (the reference search hands over elements of device
    tensors, ``batch_beam_search.py:119-128``: they are stacked on the device, never read on the host)."""
    if all(torch.is_tensor(v) for v in vals):
        return torch.stack([v.reshape(()) for v in vals]).to(device=dev, dtype=torch.int64)
    return torch.as_tensor([int(v) for v in vals], dtype=torch.int64, device=dev)


class _LazySelect:
    """Per-hypothesis state handed out by ``CTCPrefixScorer.select_state`` under the reference's search: the batch state
    it was selected from plus (hypothesis, token).  Reads like the tuple ``(r (2, Tp), log psi, T, (f_min, f_max))``."""

    __slots__ = ("scorer", "impl", "batch", "i", "tok", "value")

    def __init__(self, scorer, batch, i, tok):
        self.scorer, self.batch, self.i, self.tok, self.value = scorer, batch, i, tok, None
        self.impl = scorer.impl  # the posteriors the batch state belongs to (extend_prob installs a NEW dict)

    def materialise(self):
        if self.value is None:
            r_prev, log_psi, last, olen, _, win, f_now = self.batch
            dev = r_prev.device
            hyp = _index_vector([self.i], dev)
            tok = _index_vector([self.tok], dev)
            r_new = self.scorer._select(r_prev, last, hyp.to(torch.int32), tok.to(torch.int32), olen, win, impl=self.impl)[0]
            self.value = (r_new, log_psi[hyp[0], tok[0]], self.impl["T"], f_now)
        return self.value

    def __getitem__(self, k):
        return self.materialise()[k]

    def __len__(self):
        return 4

    def __iter__(self):
        return iter(self.materialise())


class CTCPrefixScorer(BatchPartialScorerInterface):
    """Decoder interface wrapper for the CUDA CTC prefix scorer."""

    def __init__(self, ctc: torch.nn.Module, eos: int, margin: int = 0):
        self.ctc = ctc
        self.eos = eos
        # attention-window margin of CTCPrefixScoreTH (ctc_prefix_score.py:21,56-61): with margin > 0 and attention
        # weights passed to batch_score_partial the recursion is restricted to frames around the attention peak
        self.margin = int(margin)
        self.blank = 0  # scorers/ctc.py:98
        self.impl = None  # dict with the per-utterance device buffers

    # -- shared ---------------------------------------------------------------------------------
    def _setup(self, x: torch.Tensor):
        if not x.is_cuda:
            raise RuntimeError("CTCPrefixScorer (sm_100a) needs CUDA tensors: there is no CPU fallback")
        if x.dtype == torch.float64:
            # the reference follows x.dtype (ctc_prefix_score.py:38); the kernels are fp32: refuse rather than lose
            # precision silently
            raise NotImplementedError("CTCPrefixScorer (sm_100a kernels) computes in fp32: float64 encoder output is not supported")
        lib = _lib.load()
        logp = self.ctc.log_softmax(x.unsqueeze(0))[0].to(torch.float32).contiguous()  # fp16 / bf16 inputs are widened
        T, V = logp.shape
        Tp = (T + 3) & ~3
        xT = torch.empty(V, Tp, dtype=torch.float32, device=x.device)
        r0 = torch.empty(2, Tp, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(lib.lgasr_transpose_posteriors(logp.data_ptr(), T, V, T, self.blank, xT.data_ptr(), Tp, _st(x.device)))
            _lib.check(lib.lgasr_ctc_initial_state(xT.data_ptr(), T, Tp, self.blank, r0.data_ptr(), _st(x.device)))
        self.impl = dict(lib=lib, xT=xT, r0=r0, T=T, Tp=Tp, V=V, device=x.device, dtype=x.dtype)

    def _score(self, last: torch.Tensor, olen: int, r_prev: torch.Tensor, s_prev: torch.Tensor, ids, win=(0, 0)):
        m = self.impl
        n = r_prev.shape[0]
        S = 0 if ids is None else ids.shape[1]
        scores = torch.empty(n, m["V"], dtype=torch.float32, device=m["device"])
        log_psi = torch.empty(n, m["V"], dtype=torch.float32, device=m["device"])
        ids64 = None if ids is None else ids.to(torch.int64).contiguous()
        with torch.cuda.device(m["device"]):
            _lib.check(
                m["lib"].lgasr_ctc_prefix_score_win(
                    m["xT"].data_ptr(), m["T"], m["Tp"], m["V"], self.blank, self.eos, n,
                    r_prev.data_ptr(), s_prev.data_ptr(), last.data_ptr(), olen,
                    None if ids64 is None else ids64.data_ptr(), S, scores.data_ptr(), log_psi.data_ptr(),
                    int(win[0]), int(win[1]), _st(m["device"]),
                )
            )
        return scores, log_psi

    def _select(self, r_prev: torch.Tensor, last: torch.Tensor, hyp: torch.Tensor, tok: torch.Tensor, olen: int, win=(0, 0),
                impl=None):
        m = self.impl if impl is None else impl
        k = hyp.shape[0]
        r_new = torch.empty(k, 2, m["Tp"], dtype=torch.float32, device=m["device"])
        with torch.cuda.device(m["device"]):
            _lib.check(
                m["lib"].lgasr_ctc_select_state_win(
                    m["xT"].data_ptr(), m["T"], m["Tp"], self.blank, k, r_prev.data_ptr(), last.data_ptr(),
                    hyp.data_ptr(), tok.data_ptr(), olen, r_new.data_ptr(), int(win[0]), int(win[1]), _st(m["device"]),
                )
            )
        return r_new

    # -- batch API (BatchBeamSearch) ------------------------------------------------------------
    def batch_init_state(self, x: torch.Tensor):
        self._setup(x)
        return None

    def batch_score_partial(self, y, ids, state, x, att_w=None):
        """``y (n, L)`` int64 prefixes, ``ids (n, S)`` int64 candidates or None, ``state`` list of
        per-hypothesis states (None at the first position).  Returns ``(scores (n, V), batch state)``.

        ``att_w (n, T)`` (optional, with ``margin > 0``): attention weights of the hypotheses; the frame window is
        derived from them as ``ctc_prefix_score.py:143-148`` does and carried in the state (``f_min``, ``f_max``)."""
        m = self.impl
        n = len(y)
        dev = m["device"]
        olen = len(y[0]) - 1
        f_prev = (0, 1)  # ctc_prefix_score.py:93-94
        if state is None or state[0] is None:
            r_prev = m["r0"].unsqueeze(0).expand(n, 2, m["Tp"]).contiguous()
            s_prev = torch.zeros(n, dtype=torch.float32, device=dev)
        elif (all(type(s) is _LazySelect and s.value is None and s.impl is m for s in state)
              and all(s.batch is state[0].batch for s in state)):
            # the survivors of ONE earlier call (the reference search's select_state per survivor,
            # batch_beam_search.py:347-355): one lgasr_ctc_select_state launch and one gather for all of them, the
            # (hypothesis, token) pairs never visit the host
            b_r, b_psi, b_last, b_olen, _, b_win, f_prev = state[0].batch
            hyp = _index_vector([s.i for s in state], dev)
            tok = _index_vector([s.tok for s in state], dev)
            r_prev = self._select(b_r, b_last, hyp.to(torch.int32), tok.to(torch.int32), b_olen, b_win)
            s_prev = b_psi[hyp, tok].to(torch.float32).contiguous()
            for s_, r_, p_ in zip(state, r_prev, s_prev):  # a later materialise() (streaming, tests) costs nothing
                s_.value = (r_, p_, m["T"], f_prev)
        else:
            r_prev = torch.stack([s[0] for s in state]).contiguous()
            s_prev = torch.stack([s[1] for s in state]).to(torch.float32).contiguous()
            if len(state[0]) > 3:
                f_prev = (max(s[3][0] for s in state), max(s[3][1] for s in state))
        if torch.is_tensor(y):  # (n, L) prefixes: the last tokens stay on the device (no per-hypothesis sync)
            last = y[:, -1].to(device=dev, dtype=torch.int32).contiguous()
        else:
            last = torch.as_tensor([int(yi[-1]) for yi in y], dtype=torch.int32, device=dev)
        win, f_now = (0, 0), (0, 0)
        if att_w is not None and self.margin > 0:
            T = m["T"]
            f_arg = torch.matmul(att_w.to(device=dev, dtype=torch.float32), torch.arange(T, dtype=torch.float32, device=dev))
            lo_hi = torch.stack([f_arg.min(), f_arg.max()]).cpu()  # one host read, as the reference's .cpu() calls
            f_min = max(int(lo_hi[0]), f_prev[0])
            f_max = max(int(lo_hi[1]), f_prev[1])
            win = (min(f_prev[1], max(f_min - self.margin, olen, 1)), min(f_max + self.margin, T))
            f_now = (f_min, f_max)
        scores, log_psi = self._score(last, olen, r_prev, s_prev, ids, win)
        return scores.to(m["dtype"]), (r_prev, log_psi, last, olen, "lgasr", win, f_now)

    def select_state(self, state, i, new_id=None):
        if isinstance(state, list):  # list of per-hypothesis states (batchfy / unbatchfy)
            return state[i]
        if type(state) is tuple and len(state) == 7 and state[4] == "lgasr":  # batch state of this class
            # lazy: the survivor's state is computed with all its siblings' at the next batch_score_partial, or on
            # first use as a tuple (r (2, Tp), log psi, T, (f_min, f_max))
            return _LazySelect(self, state, i, new_id)
        if type(state) is tuple and len(state) == 2 and torch.is_tensor(state[0]) and state[0].dim() > 0:
            # non-batch API: (scores over the candidate ids, states over the candidate ids)
            sc, st = state
            return sc[i], st[i]
        return None if state is None else state[i]

    # -- non-batch API (BeamSearch) -------------------------------------------------------------
    def init_state(self, x: torch.Tensor):
        self._setup(x)
        return torch.zeros((), dtype=torch.float32, device=x.device), self.impl["r0"]

    def score_partial(self, y, ids, state, x):
        """1-D prefix ``y``, 1-D candidates ``ids`` -> ``(scores (len(ids),), state)``."""
        m = self.impl
        dev = m["device"]
        prev_score, r_prev = state
        r_prev = r_prev.reshape(1, 2, m["Tp"]).contiguous()
        olen = len(y) - 1
        last = torch.as_tensor([int(y[-1])], dtype=torch.int32, device=dev)
        ids = ids.to(dev).to(torch.int64).reshape(1, -1)
        s0 = torch.zeros(1, dtype=torch.float32, device=dev)
        _, log_psi = self._score(last, olen, r_prev, s0, ids)
        psi = log_psi[0, ids[0]]
        S = ids.shape[1]
        hyp = torch.zeros(S, dtype=torch.int32, device=dev)
        r_new = self._select(r_prev, last, hyp, ids[0].to(torch.int32).contiguous(), olen)
        return (psi - prev_score).to(x.dtype), (psi, r_new)

    # -- streaming (scorers/ctc.py:128-157 -> ctc_prefix_score.py:222-269) ------------------------
    def extend_prob(self, x: torch.Tensor):
        """The utterance grew: ``x (T_new, D)`` is the encoder output so far.  Frames already held keep
        their old posteriors (``ctc_prefix_score.py:239``), the new ones are appended."""
        m = self.impl
        logp = self.ctc.log_softmax(x.unsqueeze(0))[0].to(torch.float32).contiguous()
        T_new, V = logp.shape
        if T_new <= m["T"]:
            return
        dev = m["device"]
        Tp = (T_new + 3) & ~3
        xT = torch.empty(V, Tp, dtype=torch.float32, device=dev)
        r0 = torch.empty(2, Tp, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(m["lib"].lgasr_transpose_posteriors(logp.data_ptr(), T_new, V, T_new, self.blank, xT.data_ptr(), Tp, _st(dev)))
            xT[:, : m["T"]] = m["xT"][:, : m["T"]]
            _lib.check(m["lib"].lgasr_ctc_initial_state(xT.data_ptr(), T_new, Tp, self.blank, r0.data_ptr(), _st(dev)))
        self.impl = dict(m, xT=xT, r0=r0, T=T_new, Tp=Tp)  # a new dict: lazy survivor states keep the one they came from

    def extend_state(self, state):
        """Per-hypothesis states ``(r (2, Tp_old), s, T_old)`` -> the same states over the extended frames:
        r^n = log-zero and r^b[t] = r^b[t-1] + x[t, blank] on the new frames (``ctc_prefix_score.py:246-269``)."""
        m = self.impl
        dev = m["device"]
        out = list(state)
        groups = {}
        for i, st in enumerate(state):
            if st is None or st[2] >= m["T"]:
                continue
            groups.setdefault((st[2], st[0].shape[-1]), []).append(i)
        for (T_old, Tp_old), idx in groups.items():
            r_old = torch.stack([state[i][0] for i in idx]).contiguous()
            r_new = torch.empty(len(idx), 2, m["Tp"], dtype=torch.float32, device=dev)
            with torch.cuda.device(dev):
                _lib.check(
                    m["lib"].lgasr_ctc_extend_state(
                        m["xT"].data_ptr(), T_old, m["T"], Tp_old, m["Tp"], self.blank, len(idx), r_old.data_ptr(),
                        r_new.data_ptr(), _st(dev),
                    )
                )
            for k, i in enumerate(idx):
                out[i] = (r_new[k], state[i][1], m["T"]) + tuple(state[i][3:])
        return out
