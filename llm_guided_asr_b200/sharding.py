"""Utterance-sharded decoding: one process per GPU, no data-path collective.

The reference shards decoding by splitting the key file over independent jobs
(``egs2/*/asr.sh:1555-1585`` ``inference_nj``) and concatenating + sorting the result files
(``:1603-1612``).  Utterances share no state (``espnet/nets/ctc_prefix_score.py:54,64-65,109,181,200``),
so the only exchange here is ONE final gather of ``(utt_id, tokens, score)`` records to rank 0.

The partition / merge logic is backend-agnostic (tested under ``gloo`` with world size 2 on CPU);
the per-shard decode is whatever callable the caller passes (``BatchBeamSearch.decode_batch`` on a
GPU box).
"""

from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def partition(lengths: Sequence[int], world: int) -> List[List[int]]:
    """Deal utterance indices to ``world`` shards: sort by length (longest first, index as the
    tie-break), then snake round-robin, so that every shard gets the same count (±1) and a similar
    sum of frames.  Deterministic, independent of the rank that calls it."""
    if world < 1:
        raise ValueError("world must be >= 1")
    order = sorted(range(len(lengths)), key=lambda i: (-int(lengths[i]), i))
    shards: List[List[int]] = [[] for _ in range(world)]
    for pos, i in enumerate(order):
        rnd, k = divmod(pos, world)
        shards[k if rnd % 2 == 0 else world - 1 - k].append(i)
    return shards


def batches(indices: Sequence[int], lengths: Sequence[int], batch_size: int) -> List[List[int]]:
    """Lock-step batches of one shard, length-bucketed: the shard's (already length-sorted)
    utterances are cut into consecutive groups so a batch pads to similar T."""
    if batch_size < 1:
        raise ValueError("batch_size must be >= 1")
    idx = sorted(indices, key=lambda i: (-int(lengths[i]), i))
    return [idx[s : s + batch_size] for s in range(0, len(idx), batch_size)]


Record = Tuple[int, List[int], float]  # (utterance index, best token sequence, score)


def gather_records(local: List[Record], rank: int, world: int, group=None) -> Optional[List[Record]]:
    """The final hypothesis gather: rank 0 receives every shard's records, sorted by utterance index
    (the ``cat | sort -k1`` of ``asr.sh:1603-1612``); other ranks return None."""
    if world == 1 or not dist.is_initialized():
        return sorted(local, key=lambda r: r[0])
    out = [None] * world if rank == 0 else None
    dist.gather_object(local, out, dst=0, group=group)
    if rank != 0:
        return None
    merged = [r for shard in out for r in shard]
    return sorted(merged, key=lambda r: r[0])


def decode_sharded(
    decode_fn: Callable[[List[int]], List[Tuple[List[int], float]]],
    lengths: Sequence[int],
    batch_size: int,
    rank: Optional[int] = None,
    world: Optional[int] = None,
    group=None,
) -> Optional[List[Record]]:
    """Decode ``len(lengths)`` utterances over ``world`` ranks.

    ``decode_fn(batch_indices)`` decodes one lock-step batch of this rank's utterances and returns
    ``[(tokens, score)]`` in the order of ``batch_indices`` (on a GPU box:
    ``BatchBeamSearch.decode_batch`` on the encoder outputs of those utterances, best hypothesis).
    Returns the merged, index-sorted records on rank 0, None elsewhere."""
    if rank is None:
        rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world is None:
        world = dist.get_world_size(group) if dist.is_initialized() else 1
    mine = partition(lengths, world)[rank]
    local: List[Record] = []
    for b in batches(mine, lengths, batch_size):
        res = decode_fn(b)
        if len(res) != len(b):
            raise RuntimeError(f"decode_fn returned {len(res)} results for a batch of {len(b)}")
        for i, (toks, score) in zip(b, res):
            local.append((int(i), [int(t) for t in toks], float(score)))
    return gather_records(local, rank, world, group)


def shard_summary(lengths: Sequence[int], world: int) -> Dict[str, List[int]]:
    """Counts and frame sums per shard (imbalance report of the driver)."""
    sh = partition(lengths, world)
    return {"count": [len(s) for s in sh], "frames": [sum(int(lengths[i]) for i in s) for s in sh]}


def bind_to_gpu_numa(device_index: int) -> Optional[List[int]]:
    """Pin the calling process to the host cores NVML reports as local to GPU ``device_index`` BEFORE it
    allocates pinned staging buffers, so that (first-touch) they land on the memory of the socket the GPU
    hangs off: with one process per GPU all pulling host buffers at PCIe rate, cross-socket traffic is what
    limits the host->device leg.  Returns the core list, or None when NVML / the cpuset does not allow it
    (nothing is changed then)."""
    import os

    try:
        import pynvml

        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = device_index
        if vis:
            ent = vis.split(",")[device_index].strip()
            if ent.isdigit():
                idx = int(ent)
            else:
                idx = None
                h = pynvml.nvmlDeviceGetHandleByUUID(ent)
        if idx is not None:
            h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        n = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n + 63) // 64)
        cpus = [64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if not cpus or len(cpus) == len(allowed):
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None
