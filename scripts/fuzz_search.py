#!/usr/bin/env python3
"""Randomised parity sweep (not a bench): fused search against the oracle on random small configurations
(T, V, beam, pre-beam ratio, number of full scorers, weights, pre-beam key, length ratios, primer, normalize_length),
each decoded with the default routing, with LGASR_V2=1 (k_step2) and with LGASR_NO_CLUSTER=1.
usage: fuzz_search.py [n_cases] [seed0]"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def run(n, seed0):
    import numpy as np
    import torch

    from llm_guided_asr_b200 import BatchBeamSearch, CTCPrefixScorer, LengthBonus
    from llm_guided_asr_b200.testing import synthetic as syn
    from oracle.beam_search_oracle import BeamSearchOracle, torch_scorer_fn

    dev = torch.device("cuda:0")
    bad = 0
    for case in range(n):
        rng = np.random.default_rng(seed0 + case)
        T = int(rng.integers(3, 90))
        V = int(rng.choice([8, 13, 30, 64, 101, 600, 2048]))
        beam = int(rng.integers(1, min(8, V - 2) + 1))
        ratio = float(rng.choice([1.0, 1.5, 2.5]))
        K = int(rng.integers(1, 4))
        ctc_w = float(rng.choice([0.0, 0.3, 0.5, 1.0]))
        eos = V - 1
        ntok = int(rng.integers(1, max(2, min(12, T // 2))))
        logp, table, _ = syn.peaky_case(seed0 * 7 + case, T, V, ntok, eos)
        names = ["decoder", "lm", "lm2"][:K]
        tabs = [table] + [syn.random_table(seed0 * 11 + case + k, table.shape[0], 3, V, scale=1.0) for k in range(1, K)]
        weights = {nm: float(rng.choice([0.2, 0.5, 0.7, 1.0])) for nm in names}
        weights["ctc"] = ctc_w
        if ctc_w == 1.0:
            weights = {"ctc": 1.0, **{nm: 0.0 for nm in names}}
        lb = float(rng.choice([0.0, 0.0, 0.4]))
        weights["length_bonus"] = lb
        key = str(rng.choice(["full", "full", names[0]])) if ctc_w not in (1.0,) else None
        maxr = float(rng.choice([0.0, 0.0, -float(rng.integers(1, 12)), 0.4]))
        if maxr < 0 and -maxr > T:  # the reference itself raises beyond T positions (r[start - 1], ctc_prefix_score.py:163-171)
            maxr = -float(T)
        minr = float(rng.choice([0.0, 0.0, 0.1]))
        norm = bool(rng.integers(0, 2))
        scorers = {nm: syn.PrefixTableScorer(t.to(dev)) for nm, t in zip(names, tabs)}
        scorers["ctc"] = CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos)
        scorers["length_bonus"] = LengthBonus(V)
        full = {nm: torch_scorer_fn(syn.PrefixTableScorer(t)) for nm, t in zip(names, tabs)}
        full["length_bonus"] = lambda ys, V=V: np.ones((len(ys), V), dtype=np.float32)
        full = {k: v for k, v in full.items() if weights.get(k, 0) != 0}
        desc = dict(case=seed0 + case, T=T, V=V, beam=beam, ratio=ratio, K=K, w=weights, key=key, maxr=maxr, minr=minr, norm=norm)
        try:
            bs = BatchBeamSearch(scorers=scorers, weights=weights, beam_size=beam, vocab_size=V, sos=eos, eos=eos,
                                 pre_beam_ratio=ratio, pre_beam_score_key=key, normalize_length=norm)
            nb = bs(logp.to(dev), maxr, minr)
            orc = BeamSearchOracle(full, weights, beam, V, eos, eos, pre_beam_ratio=ratio, pre_beam_score_key=key,
                                   normalize_length=norm)
            ref = orc.forward(logp.numpy(), maxr, minr)
            # hypotheses the CTC scorer gives probability zero (a blank inside the prefix: total = w_ctc * -1e10, where
            # fp32 cannot tell the candidates apart) are exact ties, ordered differently even by the reference itself
            # (torch.topk): only the count of such dead entries is compared
            # (and once such a prefix is extended, psi - s_prev = x - (-1e10) cancels the -1e10 in fp32 and leaves rounding
            # noise: more exact ties.  Prefixes with a blank inside only appear when the beam is wider than the live
            # candidates, V = 8 ... 30 here.)
            alive_o = [h for h in nb if float(h.score) > -1e9 and 0 not in h.yseq.tolist()[1:-1]]
            alive_r = [r for r in ref if float(r.score) > -1e9 and 0 not in list(r.yseq)[1:-1]]
            # scores: 1e-4 abs (the tests' tolerance) + 2e-6 relative: after 30+ positions at |score| ~ 200 one fp32 ulp is
            # 1.5e-5 and the two step kernels round their scans differently
            ok = len(alive_o) == len(alive_r) and (len(nb) == len(ref) or len(alive_o) < len(nb)) and all(
                h.yseq.tolist() == list(r.yseq) and abs(float(h.score) - float(r.score)) <= 1e-4 + 2e-6 * abs(float(r.score))
                for h, r in zip(alive_o, alive_r))
        except Exception as e:  # a refusal on one side only is a finding too
            ok = False
            desc["error"] = repr(e)[:200]
        if not ok:
            bad += 1
            print("MISMATCH", desc, flush=True)
            if os.environ.get("FUZZ_VERBOSE") and "error" not in desc:
                print("  ours  ", [(len(h.yseq), round(float(h.score), 4)) for h in nb][:12])
                print("  oracle", [(len(r.yseq), round(float(r.score), 4)) for r in ref][:12])
                for h, r in zip(nb, ref):
                    if h.yseq.tolist() != list(r.yseq) or abs(float(h.score) - float(r.score)) > 1e-4:
                        print("  first diff: ours", h.yseq.tolist(), float(h.score), {k: round(float(v), 4) for k, v in h.scores.items()})
                        print("            oracle", list(r.yseq), float(r.score), {k: round(float(v), 4) for k, v in r.scores.items()})
                        break
    print(f"{n} cases, {bad} mismatches (env V2={os.environ.get('LGASR_V2')} NO_CLUSTER={os.environ.get('LGASR_NO_CLUSTER')})", flush=True)
    return bad


def run_batched(n, seed0):
    """decode_batch over 2-6 utterances of different lengths (lock-step, ragged) and decode_stream (continuous batching)
    against one-utterance decodes of the same search object: identical n-best lists, bit for bit."""
    import numpy as np
    import torch

    from llm_guided_asr_b200 import BatchBeamSearch, CTCPrefixScorer, LengthBonus
    from llm_guided_asr_b200.testing import synthetic as syn

    dev = torch.device("cuda:0")
    bad = 0
    for case in range(n):
        rng = np.random.default_rng(seed0 + case)
        V = int(rng.choice([13, 30, 101, 600]))
        beam = int(rng.integers(1, 7))
        K = int(rng.integers(1, 3))
        eos = V - 1
        nu = int(rng.integers(2, 7))
        lens = [int(v) for v in rng.integers(4, 70, size=nu)]
        names = ["decoder", "lm"][:K]
        tabs = [syn.random_table(seed0 * 3 + case + k, max(lens) + 2, 3, V, scale=1.5) for k in range(K)]
        xs = [syn.peaky_case(seed0 * 5 + case * 7 + u, t, V, max(1, min(8, t // 3)), eos)[0].to(dev) for u, t in enumerate(lens)]
        weights = {nm: float(rng.choice([0.3, 0.7])) for nm in names}
        weights["ctc"] = float(rng.choice([0.3, 0.5]))
        weights["length_bonus"] = float(rng.choice([0.0, 0.3]))
        maxr = float(rng.choice([0.0, 0.0, -6.0]))
        scorers = {nm: syn.PrefixTableScorer(t.to(dev)) for nm, t in zip(names, tabs)}
        scorers["ctc"] = CTCPrefixScorer(ctc=syn.IdentityCTC(), eos=eos)
        scorers["length_bonus"] = LengthBonus(V)
        bs = BatchBeamSearch(scorers=scorers, weights=weights, beam_size=beam, vocab_size=V, sos=eos, eos=eos, pre_beam_score_key="full")
        desc = dict(case=seed0 + case, V=V, beam=beam, K=K, lens=lens, w=weights, maxr=maxr)
        try:
            single = [bs(x, maxr, 0.0) for x in xs]
            both = bs.decode_batch(xs, maxr, 0.0)
            stream = bs.decode_stream(xs, slots=int(rng.integers(1, nu + 1)), maxlenratio=maxr, minlenratio=0.0)
            ok = True
            for a, b, c in zip(single, both, stream):
                for other in (b, c):
                    ok = ok and len(a) == len(other) and all(
                        h.yseq.tolist() == g.yseq.tolist() and float(h.score) == float(g.score) for h, g in zip(a, other))
        except Exception as e:
            ok = False
            desc["error"] = repr(e)[:200]
        if not ok:
            bad += 1
            print("BATCH MISMATCH", desc, flush=True)
    print(f"{n} batched cases, {bad} mismatches (env V2={os.environ.get('LGASR_V2')} NO_CLUSTER={os.environ.get('LGASR_NO_CLUSTER')})", flush=True)
    return bad


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    if os.environ.get("FUZZ_CHILD"):
        sys.exit(1 if (run(n, seed0) + run_batched(max(1, n // 4), seed0)) else 0)
    rc = 0
    for env in ({}, {"LGASR_V2": "1"}, {"LGASR_NO_CLUSTER": "1"}):
        e = dict(os.environ, FUZZ_CHILD="1", **env)
        rc |= subprocess.run([sys.executable, os.path.abspath(__file__), str(n), str(seed0)], env=e).returncode
    sys.exit(rc)
