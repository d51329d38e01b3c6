"""Timestamp-rule full scorer of the OWSM (``s2t``) recipe, evaluated on the device.

Mirrors ``ScoreFilter`` of the reference (``espnet2/bin/s2t_inference.py:47-146``): same constructor
arguments, same ``score`` / ``batch_score`` contract (``(n, V)`` scores that are 0.0 or -inf, no state).  The
reference applies the rules per hypothesis in Python; here one launch of ``k_score_filter``
(``lgasr_score_filter``) reads the prefixes where they already are -- for this package's ``BatchBeamSearch``
that is the handle's own ``yseq`` array -- and writes the rows ``lgasr_step`` fuses as one more full-scorer term.
"""

import ctypes as C
from typing import Any, List, Tuple

import torch

from . import _lib
from .scorer_interface import BatchScorerInterface


class ScoreFilter(BatchScorerInterface, torch.nn.Module):
    def __init__(self, notimestamps: int, first_time: int, last_time: int, sos: int, eos: int, vocab_size: int):
        torch.nn.Module.__init__(self)
        self.notimestamps = notimestamps
        self.first_time = first_time
        self.last_time = last_time
        self.sos = sos
        self.eos = eos
        self.vocab_size = vocab_size
        # same device/dtype carrier as the reference (:73-74)
        self.param = torch.nn.Parameter(torch.tensor(0.0, dtype=torch.float32))
        self._pad = None

    def _launch(self, ys: torch.Tensor, ld_y: int, L: int, rows: int, out: torch.Tensor) -> torch.Tensor:
        if ys.device.type != "cuda":
            raise RuntimeError("ScoreFilter (sm_100a) needs CUDA tensors: there is no CPU fallback")
        lib = _lib.load()
        with torch.cuda.device(ys.device):
            st = int(torch.cuda.current_stream(ys.device).cuda_stream)
            _lib.check(
                lib.lgasr_score_filter(C.c_void_p(ys.data_ptr()), ld_y, L, rows, self.vocab_size, self.notimestamps,
                                       self.first_time, self.last_time, self.sos, self.eos, C.c_void_p(out.data_ptr()),
                                       out.stride(0) if rows > 1 else self.vocab_size, st)
            )
        return out

    def score(self, y: torch.Tensor, state: Any, x: torch.Tensor) -> Tuple[torch.Tensor, Any]:
        out, _ = self.batch_score(y.unsqueeze(0), [state], None)
        return out[0], None

    def batch_score(self, ys: torch.Tensor, states: List[Any], xs: torch.Tensor) -> Tuple[torch.Tensor, List[Any]]:
        ys = ys.to(torch.int64).contiguous()
        n, L = ys.shape
        out = torch.empty(n, self.vocab_size, dtype=torch.float32, device=ys.device)
        return self._launch(ys, L, L, n, out), [None] * n

    def batch_score_padded(self, yseq: torch.Tensor, n_run, done) -> torch.Tensor:
        """Device protocol of this package's search: ``yseq (B, W, L)`` is a view of the handle's prefix array
        (row stride = its capacity); rows of idle slots get the rule of whatever they hold, nobody reads them."""
        B, W, L = yseq.shape
        if yseq.dtype != torch.int64 or yseq.stride(2) != 1 or (B > 1 and yseq.stride(0) != W * yseq.stride(1)):
            yseq = yseq.to(torch.int64).contiguous()
        if self._pad is None or self._pad.shape[0] != B * W or self._pad.device != yseq.device:
            self._pad = torch.empty(B * W, self.vocab_size, dtype=torch.float32, device=yseq.device)
        return self._launch(yseq, yseq.stride(1) if W > 1 or B > 1 else L, L, B * W, self._pad)
