"""Shared helpers of the parity tests: golden loading and oracle wiring (test infrastructure)."""

import glob
import json
import os

import numpy as np
import torch

from llm_guided_asr_b200.testing import synthetic as syn
from oracle.beam_search_oracle import BeamSearchOracle, torch_scorer_fn

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# whole-search goldens (make_golden.py); streaming.npz (make_golden_streaming.py) and score_filter.npz
# (make_golden_score_filter.py) have their own tests
GOLDEN_NAMES = sorted(
    os.path.basename(p)[:-4]
    for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz"))
    if os.path.basename(p) not in ("streaming.npz", "score_filter.npz", "att_window.npz") and not os.path.basename(p).startswith("nb_")
)
# non-batch goldens (make_golden_nonbatch.py): reference BeamSearch + NumPy CTCPrefixScore
NONBATCH_NAMES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "nb_*.npz")))

# fp32 score tolerance of BASELINE.json's north_star: 1e-4 absolute.
SCORE_ATOL = 1e-4


class Golden:
    def __init__(self, name):
        self.z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
        self.spec = json.loads(str(self.z["spec"]))
        self.name = name

    def __getitem__(self, k):
        return self.z[k]

    def __contains__(self, k):
        return k in self.z.files

    @property
    def weights(self):
        c = self.spec
        return dict(decoder=1.0 - c["ctc"], ctc=c["ctc"], lm=c["lm"], length_bonus=c["pen"])

    def nbest(self):
        out = []
        for i in range(int(self.z["n_nbest"])):
            ln = int(self.z["nbest_len"][i])
            out.append((self.z["nbest_yseq"][i, :ln].tolist(), float(self.z["nbest_score"][i])))
        return out


def torch_scorers(g: Golden, device="cpu"):
    """The scorers dict in the order of asr_inference.py:173-210: decoder, ctc (added by the caller),
    length_bonus, lm."""
    V = g.spec["V"]
    dec = syn.PrefixTableScorer(torch.from_numpy(g["table"]).to(device))
    lm = syn.PrefixTableScorer(torch.from_numpy(g["table_lm"]).to(device))
    return dec, lm, V


def oracle_from_golden(g: Golden) -> BeamSearchOracle:
    c = g.spec
    dec, lm, V = torch_scorers(g)
    full = dict(
        decoder=torch_scorer_fn(dec),
        length_bonus=lambda ys: np.ones((len(ys), V), dtype=np.float32),
        lm=torch_scorer_fn(lm),
    )
    return BeamSearchOracle(
        full,
        g.weights,
        beam_size=c["beam"],
        vocab_size=V,
        sos=c["sos"],
        eos=c["eos"],
        pre_beam_score_key=c["key"],
        normalize_length=c.get("norm", False),
        hyp_primer=c.get("primer"),
    )
