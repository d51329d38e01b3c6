"""B200-native fused CTC/attention/LLM beam-search scorer (drop-in for ESPnet ``BatchBeamSearch``).

Importing the package is cheap and GPU-free; the CUDA library (``csrc/liblgasr.so``) is loaded
on first use by ``llm_guided_asr_b200._lib`` and its absence is a hard error (no CPU fallback).
"""

__version__ = "0.1.0"

from .beam_search import BatchBeamSearch, Hypothesis  # noqa: E402,F401
from .ctc_prefix_scorer import CTCPrefixScorer  # noqa: E402,F401
from .length_bonus import LengthBonus  # noqa: E402,F401
from .scorer_interface import (  # noqa: E402,F401
    BatchPartialScorerInterface,
    BatchScorerInterface,
    PartialScorerInterface,
    ScorerInterface,
)
