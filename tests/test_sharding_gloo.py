"""CPU: the utterance-sharded driver's host logic under gloo, world size 2 (task statement §5).
The per-shard decode is the ORACLE here (tests may use it); on a GPU box it is BatchBeamSearch."""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from llm_guided_asr_b200 import sharding
from llm_guided_asr_b200.testing import synthetic as syn
from oracle.beam_search_oracle import BeamSearchOracle, torch_scorer_fn

V, BEAM, EOS = 40, 3, 39
LENGTHS = [21, 9, 30, 14, 14, 26, 17]


def _decode_one(i):
    logp, table, _ = syn.peaky_case(100 + i, LENGTHS[i], V, 4, EOS)
    full = dict(decoder=torch_scorer_fn(syn.PrefixTableScorer(table)))
    orc = BeamSearchOracle(full, dict(decoder=0.7, ctc=0.3), BEAM, V, EOS, EOS, pre_beam_score_key="full")
    best = orc.forward(logp.numpy(), 0.0, 0.0)[0]
    return list(best.yseq), float(best.score)


def _decode_fn(batch):
    return [_decode_one(i) for i in batch]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        out = sharding.decode_sharded(_decode_fn, LENGTHS, batch_size=2)
        q.put((rank, out))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_partition_is_balanced_and_complete():
    for world in (1, 2, 3, 4, 8):
        sh = sharding.partition(LENGTHS, world)
        assert sorted(i for s in sh for i in s) == list(range(len(LENGTHS)))
        counts = [len(s) for s in sh]
        assert max(counts) - min(counts) <= 1
    s2 = sharding.shard_summary(LENGTHS, 2)
    assert abs(s2["frames"][0] - s2["frames"][1]) <= max(LENGTHS)
    assert sharding.partition([], 2) == [[], []]
    with pytest.raises(ValueError):
        sharding.partition(LENGTHS, 0)


def test_batches_are_length_bucketed():
    b = sharding.batches([0, 1, 2, 3], LENGTHS, 3)
    assert b == [[2, 0, 3], [1]]
    with pytest.raises(ValueError):
        sharding.batches([0], LENGTHS, 0)


def test_world2_gloo_equals_single_process():
    single = sharding.decode_sharded(_decode_fn, LENGTHS, batch_size=2, rank=0, world=1)
    assert [r[0] for r in single] == list(range(len(LENGTHS)))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[1] is None
    assert got[0] == single  # tokens bit-exact, scores identical (same code on every rank)


def _decode_fn_packed(batch):
    """The array path (what FusedStepEngine.gather_best_arrays feeds): deferred read-back returning a PackedBest."""
    res = [_decode_one(i) for i in batch]
    width = max(len(t) for t, _ in res) + 2  # wider than any hypothesis, like the engine's (B, maxlen) buffer
    toks = np.full((len(res), width), 12345, dtype=np.int64)
    for r, (t, _) in enumerate(res):
        toks[r, : len(t)] = t
    pb = sharding.PackedBest(toks, [len(t) for t, _ in res], [s for _, s in res])
    return lambda: pb


def _worker_packed(rank, world, port, q, lengths):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        out = sharding.decode_sharded(_decode_fn_packed, lengths, batch_size=2)
        q.put((rank, None if out is None else out.to_list()))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [len(LENGTHS), 1])
def test_world2_gloo_packed_arrays_equal_list_path(n):
    """Arrays end to end (no per-token Python objects), deferred read-back; n = 1 leaves rank 1 with an empty shard,
    which must still take part in the gather."""
    lengths = LENGTHS[:n]
    single = sharding.decode_sharded(_decode_fn, lengths, batch_size=2, rank=0, world=1)
    packed1 = sharding.decode_sharded(_decode_fn_packed, lengths, batch_size=2, rank=0, world=1)
    assert isinstance(packed1, sharding.PackedRecords) and packed1.to_list() == [(i, t, np.float32(s)) for i, t, s in single]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_packed, args=(r, 2, port, q, lengths)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[1] is None
    assert got[0] == packed1.to_list()
