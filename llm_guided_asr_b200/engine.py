"""Thin object wrapper over the C-ABI handle (``include/lgasr.h``): one handle decodes ``B``
utterances in lock-step.  PyTorch is used for device memory and streams only."""

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import _lib


def _stream_ptr(device) -> int:
    return int(torch.cuda.current_stream(device).cuda_stream)


class FusedStepEngine:
    """Owns one ``lgasr_handle``.

    Args mirror ``BeamSearch.__init__`` (``espnet/nets/beam_search.py:37-121``) plus batch extents:
    ``term_names`` are the full scorers in ``self.full_scorers`` order, ``pre_beam_key`` is "full",
    one of ``term_names`` or None.
    """

    def __init__(
        self,
        n_utts: int,
        max_frames: int,
        vocab: int,
        beam: int,
        pre_beam: int,
        term_names: Sequence[str],
        pre_beam_key: Optional[str],
        use_ctc: bool,
        ctc_weight: float,
        sos: int,
        eos: int,
        max_len_cap: int,
        primer_len: int = 1,
        end_detect: bool = True,
        blank: int = 0,
        device="cuda:0",
    ):
        self.lib = _lib.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.LgasrError("FusedStepEngine needs a CUDA device (no CPU fallback)")
        self.term_names = list(term_names)
        if pre_beam_key is None:
            key = -2
        elif pre_beam_key == "full":
            key = -1
        else:
            key = self.term_names.index(pre_beam_key)
        self.cfg = _lib.Config(
            n_utts=n_utts,
            max_frames=max_frames,
            vocab=vocab,
            beam=beam,
            pre_beam=pre_beam,
            n_terms=len(self.term_names),
            pre_beam_key=key,
            use_ctc=1 if use_ctc else 0,
            ctc_weight=float(ctc_weight),
            blank=blank,
            sos=sos,
            eos=eos,
            max_len_cap=max_len_cap,
            primer_len=primer_len,
            end_detect=1 if end_detect else 0,
        )
        self.B, self.T, self.V, self.beam = n_utts, max_frames, vocab, beam
        self.NS = len(self.term_names) + (1 if use_ctc else 0)
        self.use_ctc = bool(use_ctc)
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.lgasr_create(C.byref(self.cfg), C.byref(self._h)))
        self._i_nrun = (C.c_int32 * n_utts)()
        self._i_done = (C.c_int32 * n_utts)()
        self._i_par = (C.c_int32 * (n_utts * beam))()
        self._i_tok = (C.c_int32 * (n_utts * beam))()

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self.lib.lgasr_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- CTCPrefixScoreTH.__init__ ------------------------------------------------------------
    def set_posteriors(self, logp: Optional[torch.Tensor], xlens: Sequence[int]):
        """``logp (B, T, V)`` fp32 log-posteriors on this device (or on the host), ``xlens[B]``."""
        xl = _lib.i32_array(xlens)
        with torch.cuda.device(self.device):
            st = _stream_ptr(self.device)
            if logp is None:
                _lib.check(self.lib.lgasr_set_posteriors(self._h, None, xl, st))
                return
            assert logp.dtype == torch.float32 and logp.is_contiguous()
            assert tuple(logp.shape) == (self.B, self.T, self.V), (tuple(logp.shape), (self.B, self.T, self.V))
            if logp.is_cuda:
                _lib.check(self.lib.lgasr_set_posteriors(self._h, logp.data_ptr(), xl, st))
            else:
                _lib.check(self.lib.lgasr_set_posteriors_host(self._h, logp.data_ptr(), xl, st))

    def begin(self, maxlen: Sequence[int], minlen: Sequence[int], primer: Sequence[int]):
        with torch.cuda.device(self.device):
            _lib.check(
                self.lib.lgasr_begin(
                    self._h, _lib.i32_array(maxlen), _lib.i32_array(minlen), _lib.i64_array(primer), _stream_ptr(self.device)
                )
            )

    def refill(self, utt: int, logp: Optional[torch.Tensor], xlen: int, maxlen: int, minlen: int):
        """Continuous batching (``lgasr_refill``): slot ``utt``, whose utterance has finished, takes a new one.
        ``logp (T, V)`` fp32 log-posteriors of the new utterance on this device (``T`` = the handle's ``max_frames``;
        frames past ``xlen`` are never read)."""
        with torch.cuda.device(self.device):
            if logp is not None:
                assert logp.is_cuda and logp.dtype == torch.float32 and logp.is_contiguous()
                assert tuple(logp.shape) == (self.T, self.V), (tuple(logp.shape), (self.T, self.V))
            _lib.check(self.lib.lgasr_refill(self._h, int(utt), None if logp is None else logp.data_ptr(), int(xlen), int(maxlen),
                                             int(minlen), _stream_ptr(self.device)))

    def _terms(self, mats, consts, weights):
        ptrs, dts, ld = [], [], None
        for m in mats:
            if m is None:
                ptrs.append(0)
                dts.append(0)
                continue
            assert m.is_cuda and m.dim() == 2 and m.stride(1) == 1 and m.shape[1] == self.V
            assert m.dtype in (torch.float32, torch.bfloat16), "term matrices are fp32 or bf16"
            ml = m.stride(0) if m.shape[0] > 1 else self.V
            assert ld is None or ld == ml, "all term matrices must share one row stride"
            ld = ml
            ptrs.append(m.data_ptr())
            dts.append(1 if m.dtype == torch.bfloat16 else 0)
        return ptrs, dts, (ld if ld is not None else self.V)

    def step(self, mats: List[Optional[torch.Tensor]], consts: List[float], weights: List[float],
             norms: Optional[List[Optional[torch.Tensor]]] = None):
        """One decode position. ``mats[k]``: ``(B*beam, V)`` fp32 or bf16 device matrix (row = utt*beam+slot)
        or None for a constant term ``consts[k]``.  ``norms[k]`` (optional): ``(B*beam,)`` fp32 per-row normaliser,
        the score of an entry is ``float(mats[k][r, v]) - norms[k][r]`` (logits + logsumexp: ``lgasr_step_ex``)."""
        ptrs, dts, ld = self._terms(mats, consts, weights)
        typed = any(dts) or (norms is not None and any(n is not None for n in norms))
        with torch.cuda.device(self.device):
            if not typed:
                _lib.check(
                    self.lib.lgasr_step(
                        self._h, _lib.ptr_array(ptrs), _lib.f32_array(consts), _lib.f32_array(weights), ld, _stream_ptr(self.device)
                    )
                )
                return
            nps = []
            for k, m in enumerate(mats):
                n = None if norms is None else norms[k]
                if n is None or m is None:
                    nps.append(0)
                    continue
                assert n.is_cuda and n.dtype == torch.float32 and n.is_contiguous() and n.numel() == self.B * self.beam
                nps.append(n.data_ptr())
            _lib.check(
                self.lib.lgasr_step_ex(
                    self._h, _lib.ptr_array(ptrs), _lib.i32_array(dts), _lib.ptr_array(nps), _lib.f32_array(consts),
                    _lib.f32_array(weights), ld, _stream_ptr(self.device),
                )
            )

    def replay(self, mats_per_step, consts, weights, ld=None):
        """``mats_per_step[i][k]`` device matrices; runs len(mats_per_step) positions without host syncs."""
        n = len(mats_per_step)
        flat = []
        for row in mats_per_step:
            for m in row:
                assert m is None or m.dtype == torch.float32, "replay takes fp32 matrices (typed terms: step)"
                flat.append(0 if m is None else m.data_ptr())
        with torch.cuda.device(self.device):
            _lib.check(
                self.lib.lgasr_decode_replay(
                    self._h, n, _lib.ptr_array(flat), _lib.f32_array(consts), _lib.f32_array(weights), ld or self.V,
                    _stream_ptr(self.device),
                )
            )

    def replay_host(self, logp_host, xlens, maxlen, minlen, primer, mats_per_step_host, consts, weights, ld=None):
        """Same workload from HOST tensors (pinned for speed): copies inside the call."""
        n = len(mats_per_step_host)
        flat = []
        for row in mats_per_step_host:
            for m in row:
                flat.append(0 if m is None else m.data_ptr())
        with torch.cuda.device(self.device):
            _lib.check(
                self.lib.lgasr_decode_replay_host(
                    self._h,
                    None if logp_host is None else logp_host.data_ptr(),
                    _lib.i32_array(xlens),
                    _lib.i32_array(maxlen),
                    _lib.i32_array(minlen),
                    _lib.i64_array(primer),
                    n,
                    _lib.ptr_array(flat),
                    _lib.f32_array(consts),
                    _lib.f32_array(weights),
                    ld or self.V,
                    _stream_ptr(self.device),
                )
            )

    def gather_best_arrays(self, normalize_length: bool = False):
        """``lgasr_gather_best`` as arrays: ``tokens (B, L)`` int64, ``lens (B,)`` int32 (0 where nothing ended),
        ``scores (B,)`` float32 - no per-token Python objects (the sharded driver's result path)."""
        B = self.cfg.n_utts
        ml = self.cfg.primer_len + self.cfg.max_len_cap + 1
        toks = np.empty((B, ml), dtype=np.int64)
        lens = np.empty(B, dtype=np.int32)
        score = np.empty(B, dtype=np.float32)
        with torch.cuda.device(self.device):
            _lib.check(
                self.lib.lgasr_gather_best(
                    self._h, 1 if normalize_length else 0, ml, toks.ctypes.data_as(_lib._PI64), lens.ctypes.data_as(_lib._PI32),
                    score.ctypes.data_as(_lib._PF), _stream_ptr(self.device),
                )
            )
        return toks, lens, score

    def gather_best(self, normalize_length: bool = False):
        """Best ended hypothesis of every utterance in one call (``lgasr_gather_best``): list of
        ``(tokens, score)``; ``([], 0.0)`` where nothing ended."""
        toks, lens, score = self.gather_best_arrays(normalize_length)
        return [(toks[b, : lens[b]].tolist(), float(score[b])) for b in range(len(lens))]

    def reorder_rows(self, dst: torch.Tensor, src: torch.Tensor, row_elems: Optional[int] = None):
        """``dst[o, b*W + j, :n] = src[o, b*W + parent[b][j], :n]`` for the hypotheses that survived the last
        ``step`` (``lgasr_reorder_rows``): parents and running counts are read on the device, so this can be
        queued right behind ``step`` without a ``poll``.  ``src``/``dst``: ``(n_outer, B*W, ...)`` or
        ``(B*W, ...)`` with contiguous rows (row and outer strides may differ between the two);
        ``row_elems`` (default: a whole ``src`` row) limits the copy to the leading elements of each row."""
        BW = self.cfg.n_utts * self.cfg.beam
        if src.dim() >= 2 and src.shape[0] == BW and (dst.shape[0] == BW):
            src, dst = src.unsqueeze(0), dst.unsqueeze(0)
        if src.shape[1] != BW or dst.shape[1] != BW or src.shape[0] != dst.shape[0] or src.dtype != dst.dtype:
            raise ValueError("reorder_rows: (n_outer, B*beam, ...) tensors of one dtype expected")
        es = src.element_size()

        def rows_contiguous(t):
            want = 1
            for k in range(t.dim() - 1, 1, -1):
                if t.shape[k] != 1 and t.stride(k) != want:
                    return False
                want *= t.shape[k]
            return t.stride(1) >= want or t.shape[1] == 1

        if not rows_contiguous(src) or not rows_contiguous(dst):
            raise ValueError("reorder_rows: the trailing dimensions of a row must be contiguous")
        n_src = int(np.prod(src.shape[2:])) if src.dim() > 2 else 1
        n_dst = int(np.prod(dst.shape[2:])) if dst.dim() > 2 else 1
        n = n_src if row_elems is None else int(row_elems)
        if n > n_src or n > n_dst:
            raise ValueError("reorder_rows: row_elems exceeds a row")
        with torch.cuda.device(self.device):
            _lib.check(
                self.lib.lgasr_reorder_rows(
                    self._h, C.c_void_p(dst.data_ptr()), C.c_void_p(src.data_ptr()), src.shape[0], dst.stride(0) * es,
                    src.stride(0) * es, dst.stride(1) * es, src.stride(1) * es, n * es, _stream_ptr(self.device),
                )
            )

    def launch_info(self) -> dict:
        """Geometry of the fused step kernel of this handle (``lgasr_launch_info``)."""
        v = [C.c_int32() for _ in range(4)]
        _lib.check(self.lib.lgasr_launch_info(self._h, *[C.byref(x) for x in v]))
        return dict(cluster=bool(v[0].value), smem_bytes=v[1].value, chunk_floats=v[2].value, stages=v[3].value)

    def set_profiling(self, on: bool):
        _lib.check(self.lib.lgasr_set_profiling(self._h, 1 if on else 0))

    def get_profile(self):
        """(ms in k_score_rows, ms in k_select, steps) accumulated since set_profiling(True)."""
        a, b, n = C.c_float(), C.c_float(), C.c_int32()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.lgasr_get_profile(self._h, C.byref(a), C.byref(b), C.byref(n)))
        return float(a.value), float(b.value), int(n.value)

    def poll(self, with_parents=True):
        """(n_running[B], done[B], parent[B][beam], token[B][beam], n_active) as numpy / int."""
        act = C.c_int32()
        with torch.cuda.device(self.device):
            _lib.check(
                self.lib.lgasr_poll(
                    self._h,
                    self._i_nrun,
                    self._i_done,
                    self._i_par if with_parents else None,
                    self._i_tok if with_parents else None,
                    C.byref(act),
                    _stream_ptr(self.device),
                )
            )
        nrun = np.ctypeslib.as_array(self._i_nrun).copy()
        done = np.ctypeslib.as_array(self._i_done).copy()
        par = tok = None
        if with_parents:
            par = np.ctypeslib.as_array(self._i_par).reshape(self.B, self.beam).copy()
            tok = np.ctypeslib.as_array(self._i_tok).reshape(self.B, self.beam).copy()
        return nrun, done, par, tok, int(act.value)

    def view(self, which: int) -> torch.Tensor:
        """Zero-copy device tensors: 0 yseq (B, beam, Lcap) int64, 1 n_running, 2 parent, 3 token,
        4 done, 5 score, 6 cur_len."""
        p, ld = C.c_void_p(), C.c_int64()
        _lib.check(self.lib.lgasr_view(self._h, which, C.byref(p), C.byref(ld)))
        B, W = self.B, self.beam
        shape, ts = {
            0: ((B, W, int(ld.value)), "<i8"),
            1: ((B,), "<i4"),
            2: ((B, W), "<i4"),
            3: ((B, W), "<i4"),
            4: ((B,), "<i4"),
            5: ((B, W), "<f4"),
            6: ((B,), "<i4"),
        }[which]
        return torch.as_tensor(_lib.DeviceView(p.value, shape, ts), device=self.device)

    def gather_ended(self, utt: int):
        """All ended hypotheses of ``utt`` in ending order: (tokens list[list[int]], score (n,), scores (n, NS))."""
        n, ml = C.c_int32(), C.c_int32()
        with torch.cuda.device(self.device):
            st = _stream_ptr(self.device)
            _lib.check(self.lib.lgasr_gather_ended(self._h, utt, C.byref(n), C.byref(ml), None, None, None, None, st))
            n, ml = int(n.value), int(ml.value)
            if n == 0:
                return [], np.zeros((0,), np.float32), np.zeros((0, self.NS), np.float32)
            toks = np.empty((n, ml), np.int64)
            lens = np.empty((n,), np.int32)
            score = np.empty((n,), np.float32)
            scores = np.empty((n, max(self.NS, 1)), np.float32)
            _lib.check(
                self.lib.lgasr_gather_ended(
                    self._h,
                    utt,
                    None,
                    None,
                    toks.ctypes.data_as(_lib._PI64),
                    lens.ctypes.data_as(_lib._PI32),
                    score.ctypes.data_as(_lib._PF),
                    scores.ctypes.data_as(_lib._PF),
                    st,
                )
            )
        return [toks[i, : lens[i]].tolist() for i in range(n)], score, scores[:, : self.NS]
