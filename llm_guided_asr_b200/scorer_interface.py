"""Scorer plug-in API of the beam search (the drop-in boundary).

Mirrors the four duck-typed interfaces of the reference,
``espnet/nets/scorer_interface.py:9-186`` (ScorerInterface :9, BatchScorerInterface :83,
PartialScorerInterface :129, BatchPartialScorerInterface :162): same method names,
argument meaning and default behaviour, so a scorer written against either set works
with either search.

The reference search classifies scorers with ``isinstance`` (``espnet/nets/beam_search.py:86-95``).
When the reference package is importable (``PYTHONPATH=/root/reference`` in the build
container) its classes are re-exported here, so that scorers from this package plug into
the *unmodified* reference ``BatchBeamSearch`` and vice versa.  On a box without the
reference the local definitions below are used.
"""

from typing import Any, List, Tuple

import torch

try:  # pragma: no cover - depends on the environment
    from espnet.nets.scorer_interface import (  # type: ignore
        BatchPartialScorerInterface,
        BatchScorerInterface,
        PartialScorerInterface,
        ScorerInterface,
    )

    USING_REFERENCE_INTERFACES = True
except Exception:  # reference not on the path: local, API-identical definitions
    USING_REFERENCE_INTERFACES = False

    class ScorerInterface:
        """Scores every token of the vocabulary for one hypothesis."""

        def init_state(self, x: torch.Tensor) -> Any:
            """Initial scorer state for the encoded utterance ``x`` (optional)."""
            return None

        def select_state(self, state: Any, i: int, new_id: int = None) -> Any:
            """State of hypothesis ``i`` picked out of a batched state (optional)."""
            return None if state is None else state[i]

        def score(self, y: torch.Tensor, state: Any, x: torch.Tensor) -> Tuple[torch.Tensor, Any]:
            """``(n_vocab,)`` next-token scores for the int64 prefix ``y`` (required)."""
            raise NotImplementedError

        def final_score(self, state: Any) -> float:
            """Score added when the hypothesis ends (optional)."""
            return 0.0

    class BatchScorerInterface(ScorerInterface):
        """Scores every token for a batch of hypotheses."""

        def batch_init_state(self, x: torch.Tensor) -> Any:
            return self.init_state(x)

        def batch_score(
            self, ys: torch.Tensor, states: List[Any], xs: torch.Tensor
        ) -> Tuple[torch.Tensor, List[Any]]:
            """``(n, n_vocab)`` scores for int64 prefixes ``ys (n, L)``; default loops over ``score``."""
            rows, new_states = [], []
            for y, st, x in zip(ys, states, xs):
                r, s = self.score(y, st, x)
                rows.append(r)
                new_states.append(s)
            return torch.cat(rows, 0).view(ys.shape[0], -1), new_states

    class PartialScorerInterface(ScorerInterface):
        """Scores only a pre-pruned set of next tokens (e.g. CTC prefix scoring)."""

        def score_partial(
            self, y: torch.Tensor, next_tokens: torch.Tensor, state: Any, x: torch.Tensor
        ) -> Tuple[torch.Tensor, Any]:
            raise NotImplementedError

    class BatchPartialScorerInterface(BatchScorerInterface, PartialScorerInterface):
        """Batched partial scorer; returns a full ``(n, n_vocab)`` matrix."""

        def batch_score_partial(
            self, ys: torch.Tensor, next_tokens: torch.Tensor, states: List[Any], xs: torch.Tensor
        ) -> Tuple[torch.Tensor, Any]:
            raise NotImplementedError


__all__ = [
    "ScorerInterface",
    "BatchScorerInterface",
    "PartialScorerInterface",
    "BatchPartialScorerInterface",
    "USING_REFERENCE_INTERFACES",
]
