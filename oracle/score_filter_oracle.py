"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the OWSM timestamp-rule scorer.

Follows ``ScoreFilter.score`` of the reference (``espnet2/bin/s2t_inference.py:47-118``): a full scorer whose
``(n, V)`` output is 0.0 or -inf.  Pinned to ``tests/golden/score_filter.npz``, recorded from the reference's
own class (``tests/golden/make_golden_score_filter.py``).  Only ``tests/`` may import this module.
"""

import numpy as np


def score_filter_mask(y, V, sos, eos, notimestamps, first_time, last_time):
    """Boolean mask (V,) of the entries the reference sets to -inf for the int prefix ``y``."""
    y = np.asarray(y, dtype=np.int64)
    neg = np.zeros(V, dtype=bool)
    if np.any(y == notimestamps):  # :93-95  no time prediction: timestamps are suppressed
        neg[first_time : last_time + 1] = True
    elif len(y) >= 3 and y[-3] == sos:  # :96-99  the first token must be a timestamp
        neg[:first_time] = True
        neg[last_time + 1 :] = True
    else:
        times = y[(y >= first_time) & (y <= last_time)]  # :101
        if len(times) % 2 == 1:  # :102-106  an open segment: no eos, timestamps do not go back
            neg[eos] = True
            neg[first_time : times[-1] + 1] = True
        elif first_time <= y[-1] <= last_time:  # :109-113  segment just closed: timestamp (not earlier) or eos
            neg[: y[-1]] = True
            neg[last_time + 1 :] = True
            neg[eos] = False
        else:  # :114-116  illegal hypothesis
            neg[:] = True
    return neg


def score_filter(ys, V, sos, eos, notimestamps, first_time, last_time):
    """``(n, V)`` float32 scores (0.0 / -inf) for the rows of ``ys (n, L)``."""
    out = np.zeros((len(ys), V), dtype=np.float32)
    for i, y in enumerate(ys):
        out[i, score_filter_mask(y, V, sos, eos, notimestamps, first_time, last_time)] = -np.inf
    return out
