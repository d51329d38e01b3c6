"""Length bonus scorer (API mirror of ``espnet/nets/scorers/length_bonus.py:10-60``).

Scores every token +1; the fused step folds it in as a scalar term (``constant_score``) instead of
reading an ``(n, V)`` matrix of ones."""

from typing import Any, List, Tuple

import torch

from .scorer_interface import BatchScorerInterface


class LengthBonus(BatchScorerInterface):
    constant_score = 1.0

    def __init__(self, n_vocab: int):
        self.n = n_vocab

    def score(self, y, state, x):
        return torch.tensor([1.0], device=x.device, dtype=x.dtype).expand(self.n), None

    def batch_score(self, ys: torch.Tensor, states: List[Any], xs: torch.Tensor) -> Tuple[torch.Tensor, List[Any]]:
        return torch.tensor([1.0], device=xs.device, dtype=xs.dtype).expand(ys.shape[0], self.n), None
