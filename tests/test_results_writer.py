"""CPU: result files in the layout of the reference's decoding stage (asr_inference.py:653-676, 895-907)."""

import os

import pytest
import torch

from llm_guided_asr_b200.beam_search import Hypothesis
from llm_guided_asr_b200.results import NBestWriter, hypothesis_to_result


def _hyp(tokens, score):
    return Hypothesis(yseq=torch.tensor(tokens, dtype=torch.int64), score=torch.tensor(score), scores={}, states={})


def test_hypothesis_to_result_drops_sos_eos_and_blank():
    tl = ["<blank>", "<unk>", "a", "b", "c", "<sos/eos>"]
    text, token, token_int, hyp = hypothesis_to_result(_hyp([5, 2, 0, 3, 4, 5], -1.5), tl, lambda t: "".join(t))
    assert token_int == [2, 3, 4] and token == ["a", "b", "c"] and text == "abc"
    assert hypothesis_to_result(_hyp([5, 5], 0.0), tl)[2] == []


def test_writer_layout(tmp_path):
    tl = ["<blank>", "<unk>", "a", "b", "c", "<sos/eos>"]
    with NBestWriter(str(tmp_path), tl, lambda t: " ".join(t), nbest=2) as w:
        w.write("utt1", [_hyp([5, 2, 3, 5], -1.0), _hyp([5, 4, 5], -2.5), _hyp([5, 5], -9.0)])
        w.write("utt2", [_hyp([5, 4, 4, 5], -0.25)])
        with pytest.raises(RuntimeError):
            w.write("utt1", [_hyp([5, 5], 0.0)])
    rd = lambda n, name: open(os.path.join(tmp_path, f"{n}best_recog", name)).read().splitlines()  # noqa: E731
    assert rd(1, "token") == ["utt1 a b", "utt2 c c"]
    assert rd(1, "token_int") == ["utt1 2 3", "utt2 4 4"]
    assert rd(1, "score") == ["utt1 -1.0", "utt2 -0.25"]
    assert rd(1, "text") == ["utt1 a b", "utt2 c c"]
    assert rd(2, "token_int") == ["utt1 4"]
    assert not os.path.exists(os.path.join(tmp_path, "3best_recog"))
