"""GPU: randomised parity sweep of the fused search against the oracle (scripts/fuzz_search.py: random T, V, beam,
pre-beam ratio, 1-3 full scorers, weights incl. CTC-only and no-CTC, pre-beam key, length ratios, normalize_length), under
the default kernel routing, with k_step2 forced (LGASR_V2=1) and with the ticket exchange (LGASR_NO_CLUSTER=1).
Bar: tokens bit-exact, scores within 1e-4 abs; a refusal on one side only counts as a mismatch."""

import os
import sys

import pytest

pytestmark = pytest.mark.gpu

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))


@pytest.mark.parametrize("env", [{}, {"LGASR_V2": "1"}, {"LGASR_NO_CLUSTER": "1"}, {"LGASR_V2": "1", "LGASR_DENSE": "1"}])
def test_random_configurations_match_the_oracle(monkeypatch, env):
    import fuzz_search

    for k, v in env.items():
        monkeypatch.setenv(k, v)
    assert fuzz_search.run(40, 5000) == 0
    assert fuzz_search.run_batched(10, 6000) == 0  # decode_batch / decode_stream == one-utterance decodes, bit for bit
