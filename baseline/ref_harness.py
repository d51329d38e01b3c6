"""Harness around the UNMODIFIED reference (``baseline/_ref``, installed by ``scripts/install_ref.sh``).

Only wiring lives here, no arithmetic: the scorer / weight set-up of ``espnet2/bin/asr_inference.py:173-181,309-315,
352-362,379`` restated around the reference's own classes (``espnet2.bin.asr_inference`` itself does not import in this
image: ``hydra``, ``humanfriendly`` are absent).  Used by ``bench.py --impl reference`` (CPU),
``bench.py --impl reference_gpu`` (the reference's GPU PyTorch path, the north_star's 20x denominator) and by the
"scorer only" adoption tests.  Nothing of ``llm_guided_asr_b200`` is imported here.
"""

import os
import sys
import time

REF_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_DIR, "espnet", "nets"))


def activate() -> None:
    """Put the installed reference first on ``sys.path`` (idempotent)."""
    if not available():
        raise RuntimeError(f"reference not installed: run scripts/install_ref.sh (expected {REF_DIR}/espnet)")
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    sys.dont_write_bytecode = True


def build_search(tables, cfg, device="cpu", lm_tables=None, lm_weight=0.0, row0=0, ctc_scorer_cls=None):
    """Reference ``BatchBeamSearch`` over table-driven full scorers.

    ``tables (L, rows, V)``: the decoder's log-prob matrix of every position (row = running-hypothesis slot, the
    bench workload of BASELINE.md §3); ``cfg``: V, beam, ctc_weight.  Scorers dict in the order of
    ``asr_inference.py:173-210``: decoder, ctc, length_bonus (weight 0 -> dropped, ``beam_search.py:82-85``), lm."""
    activate()
    import torch
    from espnet.nets.batch_beam_search import BatchBeamSearch
    from espnet.nets.beam_search import BeamSearch
    from espnet.nets.scorer_interface import BatchScorerInterface
    from espnet.nets.scorers.ctc import CTCPrefixScorer

    if ctc_scorer_cls is not None:  # "scorer only" adoption: the caller's drop-in class under the reference search
        CTCPrefixScorer = ctc_scorer_cls
    from espnet.nets.scorers.length_bonus import LengthBonus

    class SlotTable(BatchScorerInterface):
        """``batch_score`` returns ``tables[position, row0 : row0 + n]`` (bit-identical on every device)."""

        def __init__(self, t):
            self.t = t

        def score(self, y, state, x):
            return self.t[y.shape[0] - 1, row0], None

        def batch_score(self, ys, states, xs):
            n = ys.shape[0]
            return self.t[ys.shape[1] - 1, row0 : row0 + n], [None] * n

    class IdentityCTC(torch.nn.Module):
        """The "encoder output" already is the (T, V) log-posterior matrix (stand-in for ``espnet2/asr/ctc.py:196-204``)."""

        def log_softmax(self, hs_pad):
            return hs_pad.clone()

    V, beam = cfg["V"], cfg["beam"]
    eos = V - 1
    scorers = dict(
        decoder=SlotTable(tables.to(device)),
        ctc=CTCPrefixScorer(ctc=IdentityCTC(), eos=eos),
        length_bonus=LengthBonus(V),
    )
    weights = dict(decoder=1.0 - cfg["ctc_weight"], ctc=cfg["ctc_weight"], lm=lm_weight, ngram=0.0, length_bonus=0.0)
    if lm_tables is not None:
        scorers["lm"] = SlotTable(lm_tables.to(device))
    bs = BeamSearch(
        beam_size=beam, weights=weights, scorers=scorers, sos=eos, eos=eos, vocab_size=V, token_list=None,
        pre_beam_score_key=None if cfg["ctc_weight"] == 1.0 else "full",
    )
    bs.__class__ = BatchBeamSearch  # asr_inference.py:379
    bs.to(device=device, dtype=torch.float32).eval()
    return bs


def table_parts(tables, device="cpu", row0=0):
    """The two stand-ins of ``build_search`` on their own (for a search object built by the caller): a table-driven full
    scorer (``BatchScorerInterface`` of the reference) and the identity CTC head."""
    activate()
    import torch
    from espnet.nets.scorer_interface import BatchScorerInterface

    t = tables.to(device)

    class SlotTable(BatchScorerInterface):
        def score(self, y, state, x):
            return t[y.shape[0] - 1, row0], None

        def batch_score(self, ys, states, xs):
            n = ys.shape[0]
            return t[ys.shape[1] - 1, row0 : row0 + n], [None] * n

    class IdentityCTC(torch.nn.Module):
        def log_softmax(self, hs_pad):
            return hs_pad.clone()

    return SlotTable(), IdentityCTC()


def decode_one(logp_b, tables_b, cfg, positions, device="cpu", lm_tables_b=None, lm_weight=0.0, ctc_scorer_cls=None):
    """One utterance through the reference search.  Returns ``(seconds, [(tokens, score), ...] n-best)``; the clock
    runs around ``beam_search(x=...)`` only (``time.perf_counter``, with ``cuda.synchronize`` on both sides on a GPU)."""
    import torch

    bs = build_search(tables_b, cfg, device, lm_tables_b, lm_weight, ctc_scorer_cls=ctc_scorer_cls)
    x = logp_b.to(device)
    with torch.no_grad():
        if str(device).startswith("cuda"):
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        nb = bs(x=x, maxlenratio=-float(positions), minlenratio=0.0)
        if str(device).startswith("cuda"):
            torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    return dt, [([int(t) for t in h.yseq.tolist()], float(h.score)) for h in nb]


_SHARED = {}


def _worker(b):
    """One utterance in a pool worker.  The inputs are inherited through fork (``_SHARED``), only the index travels."""
    import torch

    torch.set_num_threads(1)
    sh = _SHARED
    W = sh["cfg"]["beam"]
    L = sh["positions"]
    return decode_one(sh["logp"][b], sh["tables"][:L, b * W : (b + 1) * W], sh["cfg"], L)


class CpuPool:
    """``procs`` single-threaded reference processes (job-level sharding as ``egs2/*/asr.sh`` ``inference_nj``), forked
    once with the inputs in their address space; ``run(utts)`` decodes the utterances and returns (wall seconds of the
    decodes only - no process start-up, no input pickling -, results)."""

    def __init__(self, logp, tables, cfg, positions, procs):
        import multiprocessing as mp

        _SHARED.update(logp=logp, tables=tables, cfg=cfg, positions=positions)
        self.procs = procs
        self.pool = mp.get_context("fork").Pool(procs) if procs > 1 else None
        if self.pool is not None:  # start the workers' interpreters before the clock does
            self.pool.map(int, range(procs), chunksize=1)

    def run(self, utts):
        t0 = time.perf_counter()
        res = [_worker(b) for b in utts] if self.pool is None else self.pool.map(_worker, list(utts), chunksize=1)
        return time.perf_counter() - t0, [r[1] for r in res]

    def close(self):
        if self.pool is not None:
            self.pool.close()
            self.pool.join()
            self.pool = None


def decode_many_cpu(logp, tables, cfg, utts, positions, procs):
    """``utts`` utterance indices over ``procs`` reference processes.  ``tables (positions, B*beam, V)``.
    Returns (wall seconds, results)."""
    pool = CpuPool(logp, tables, cfg, positions, procs)
    try:
        return pool.run(utts)
    finally:
        pool.close()
