"""Record goldens for the timestamp-rule scorer of the OWSM recipe from the REFERENCE's own class.

``espnet2.bin.s2t_inference`` cannot be imported in this container (it pulls in ``hydra``), so the
``ScoreFilter`` class definition is taken out of the reference file with ``ast`` and executed as it stands
(nothing is copied into this repository).  Run in the build container only:

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_score_filter.py

Output: tests/golden/score_filter.npz -- prefixes (padded, with lengths) and, per prefix, the mask of
vocabulary entries the reference sets to -inf (everything else is 0.0).
"""

import ast
import os
from typing import Any, List, Tuple  # noqa: F401  (names the class body refers to)

import numpy as np
import torch

REF = "/root/reference/espnet2/bin/s2t_inference.py"
HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference_class():
    src = open(REF).read()
    tree = ast.parse(src)
    node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "ScoreFilter")
    code = ast.get_source_segment(src, node)

    class BatchScorerInterface:  # the base only marks the type (espnet/nets/scorer_interface.py:83)
        pass

    ns = dict(torch=torch, np=np, Any=Any, List=List, Tuple=Tuple, BatchScorerInterface=BatchScorerInterface)
    exec(compile(code, REF, "exec"), ns)
    return ns["ScoreFilter"]


def main():
    cls = load_reference_class()
    V, sos, eos, notime, first, last = 90, 1, 2, 3, 40, 79
    lang, task = 5, 6
    f = cls(notimestamps=notime, first_time=first, last_time=last, sos=sos, eos=eos, vocab_size=V)
    rng = np.random.RandomState(0)
    prefixes = [
        [sos, lang, notime],  # no timestamps predicted
        [sos, lang, notime, 10, 11],
        [sos, lang, task],  # y[-3] == sos: first token must be a timestamp
        [sos, lang, task, 40],  # odd number of timestamps
        [sos, lang, task, 45, 12, 13],
        [sos, lang, task, 45, 12, 50],  # even, last is a timestamp
        [sos, lang, task, 45, 12, 79],
        [sos, lang, task, 45, 12, 50, 50],  # odd again, monotone bound at 50
        [sos, lang, task, 45, 12, 50, 14],  # even, last is text: illegal
        [sos, lang, task, 79],
        [sos, lang, task, 41, 7, 41, 41, 8, 8, 60],
    ]
    for _ in range(40):  # random walks over text / timestamp tokens
        n = rng.randint(1, 12)
        body = [int(rng.randint(first, last + 1)) if rng.rand() < 0.45 else int(rng.randint(7, first)) for _ in range(n)]
        prefixes.append([sos, lang, task if rng.rand() < 0.8 else notime] + body)
    L = max(len(p) for p in prefixes)
    ys = np.zeros((len(prefixes), L), dtype=np.int64)
    lens = np.zeros(len(prefixes), dtype=np.int32)
    masks = np.zeros((len(prefixes), V), dtype=bool)
    for i, p in enumerate(prefixes):
        ys[i, : len(p)] = p
        lens[i] = len(p)
        s, _ = f.score(torch.tensor(p, dtype=torch.int64), None, None)
        s = s.detach().numpy()
        assert np.all((s == 0) | np.isneginf(s))
        masks[i] = np.isneginf(s)
    # the batch entry point agrees with the per-prefix one on equal-length prefixes
    same = [i for i in range(len(prefixes)) if lens[i] == 6]
    bs, _ = f.batch_score(torch.tensor(ys[same, :6]), [None] * len(same), [None] * len(same))
    assert np.array_equal(np.isneginf(bs.detach().numpy()), masks[same])
    np.savez_compressed(os.path.join(HERE, "score_filter.npz"), ys=ys, lens=lens, masks=masks,
                        spec=np.array([V, sos, eos, notime, first, last], dtype=np.int64))
    print("wrote", len(prefixes), "prefixes; -inf fraction", masks.mean())


if __name__ == "__main__":
    main()
