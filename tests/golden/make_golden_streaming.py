#!/usr/bin/env python3
"""Golden vectors of the streaming extension (CTCPrefixScoreTH.extend_prob / extend_state,
espnet/nets/ctc_prefix_score.py:222-269) recorded from the UNMODIFIED reference, imported from
/root/reference (build container only):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_streaming.py

Scenario (one utterance, blank 0): score the first position on T1 frames, keep two (hyp, token)
survivors with select_state, extend the posteriors to T2 frames and the survivor states with
extend_state, score the second position on T2 frames; then once more to T3."""

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference")
sys.dont_write_bytecode = True

from espnet.nets.ctc_prefix_score import CTCPrefixScoreTH  # noqa: E402

V, EOS = 40, 39
T1, T2, T3 = 12, 20, 31
g = torch.Generator().manual_seed(91)
logp = torch.log_softmax(3.0 * torch.randn(T3, V, generator=g), dim=-1)
# the encoder of a streaming model re-emits earlier frames with slightly different values: the
# reference keeps the OLD posteriors of frames it already holds (:239)
logp2 = logp[:T2] + 0.01 * torch.randn(T2, V, generator=g)
logp3 = logp[:T3] + 0.01 * torch.randn(T3, V, generator=g)

out = dict(V=V, eos=EOS, T=np.array([T1, T2, T3]), logp1=logp[:T1].numpy(), logp2=logp2.numpy(), logp3=logp3.numpy())
impl = CTCPrefixScoreTH(logp[:T1].clone().unsqueeze(0), torch.tensor([T1]), 0, EOS)

# position 0: one hypothesis [sos]
y = [torch.tensor([EOS])]
ids0 = torch.tensor([[3, 7, 11, 20, 25, EOS]])
sc0, st0 = impl(y, None, ids0)
out.update(ids0=ids0.numpy(), sc0=sc0[0, ids0[0]].numpy())
r, log_psi, f_min, f_max, idmap = st0
surv = [(0, 7), (0, 20)]
states = []
for h, tok in surv:  # CTCPrefixScorer.select_state, TH branch (scorers/ctc.py:56-62)
    states.append((r[:, :, h, idmap[h, tok]], log_psi[h, tok].expand(log_psi.size(1)), f_min, f_max))
out.update(surv=np.array(surv), r_sel=np.stack([s[0].numpy() for s in states]), s_sel=np.array([float(s[1][0]) for s in states]))

for stage, (lp, T) in enumerate(((logp2, T2), (logp3, T3)), start=1):
    impl.extend_prob(lp.clone().unsqueeze(0))
    states = [impl.extend_state(s) for s in states]
    out[f"r_ext{stage}"] = np.stack([s[0].numpy() for s in states])
    ys = [torch.tensor([EOS, 7] + [5] * (stage - 1)), torch.tensor([EOS, 20] + [20] * (stage - 1))]
    ids = torch.tensor([[5, 7, 9, 20, 30, EOS], [2, 20, 21, 22, 33, 38]])
    batch_state = (torch.stack([s[0] for s in states], dim=2), torch.stack([s[1] for s in states]), f_min, f_max)
    sc, st = impl(ys, batch_state, ids)
    out[f"ids{stage}"] = ids.numpy()
    out[f"sc{stage}"] = torch.gather(sc, 1, ids).numpy()
    out[f"sc{stage}_eos"] = sc[:, EOS].numpy()
    r, log_psi, f_min, f_max, idmap = st
    nxt = [(0, 5), (1, 20)]
    states = [(r[:, :, h, idmap[h, tok]], log_psi[h, tok].expand(log_psi.size(1)), f_min, f_max) for h, tok in nxt]
    out[f"r_sel{stage}"] = np.stack([s[0].numpy() for s in states])
    out[f"s_sel{stage}"] = np.array([float(s[1][0]) for s in states])

np.savez_compressed(os.path.join(HERE, "streaming.npz"), **out)
print("wrote streaming.npz", {k: np.asarray(v).shape for k, v in out.items()})
