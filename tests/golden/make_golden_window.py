#!/usr/bin/env python3
"""Goldens of the attention-window path of the reference's ``CTCPrefixScoreTH`` (``margin > 0`` with ``att_w``,
``espnet/nets/ctc_prefix_score.py:56-61,143-148``; used by the ESPnet1 decoders, ``rnn/decoders.py:665,824-826``).
A scripted sequence of calls: synthetic attention weights peaked on a frame that moves forward, the survivor of every
call chosen by a fixed rule.  Run in the build container only:

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_window.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, "/root/reference")
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True

from espnet.nets.ctc_prefix_score import CTCPrefixScoreTH  # noqa: E402

from llm_guided_asr_b200.testing import synthetic as syn  # noqa: E402


def main():
    T, V, n, S, margin, ncalls = 48, 40, 3, 5, 4, 6
    eos = V - 1
    logp = syn.random_logp(77, T, V)
    impl = CTCPrefixScoreTH(logp.clone().unsqueeze(0), torch.tensor([T]), 0, eos, margin=margin)
    rng = np.random.default_rng(5)
    out = dict(logp=logp.numpy(), T=np.array(T), V=np.array(V), n=np.array(n), S=np.array(S), margin=np.array(margin),
               eos=np.array(eos), ncalls=np.array(ncalls))
    y = [torch.tensor([eos]) for _ in range(n)]
    state = None
    for c in range(ncalls):
        centre = 3.4 + 7 * c  # off the integers: int(f_arg) must not hinge on the last bit of a matmul
        fr = torch.arange(T, dtype=torch.float32)
        att = torch.stack([torch.softmax(-0.5 * (fr - (centre + 1.1 * h)) ** 2, dim=0) for h in range(n)])  # (n, T)
        ids = torch.from_numpy(np.stack([rng.choice(np.arange(1, V - 1), size=S, replace=False) for _ in range(n)]))
        if c == 2:  # a repeated token among the candidates
            ids[1, 0] = int(y[1][-1])
        scores, new_state = impl(y, state, ids, att)
        r, log_psi, f_min, f_max, idmap = new_state
        out[f"c{c}_att"] = att.numpy()
        out[f"c{c}_ids"] = ids.numpy()
        out[f"c{c}_last"] = np.array([int(v[-1]) for v in y])
        out[f"c{c}_olen"] = np.array(len(y[0]) - 1)
        out[f"c{c}_scores_ids"] = torch.gather(scores, 1, ids).numpy()
        out[f"c{c}_scores_eos"] = scores[:, eos].numpy()
        out[f"c{c}_f"] = np.array([f_min, f_max])
        # survivors: hypothesis h keeps candidate (c + h) % S of hypothesis (h + 1) % n
        hyp = [(h + 1) % n for h in range(n)]
        tok = [int(ids[hyp[h], (c + h) % S]) for h in range(n)]
        out[f"c{c}_sel_hyp"] = np.array(hyp)
        out[f"c{c}_sel_tok"] = np.array(tok)
        r_new = torch.stack([r[:, :, hyp[h], idmap[hyp[h], tok[h]]] for h in range(n)], dim=2)  # (T, 2, n)
        s_new = torch.stack([log_psi[hyp[h], tok[h]] for h in range(n)])
        out[f"c{c}_r_new"] = r_new.numpy()
        out[f"c{c}_s_new"] = s_new.numpy()
        y = [torch.cat([y[hyp[h]], torch.tensor([tok[h]])]) for h in range(n)]
        state = (r_new, s_new.unsqueeze(1).expand(n, V).contiguous(), f_min, f_max)
    np.savez_compressed(os.path.join(HERE, "att_window.npz"), **out)
    print("att_window.npz", {k: out[k].tolist() for k in out if k.endswith("_f")})


if __name__ == "__main__":
    main()
