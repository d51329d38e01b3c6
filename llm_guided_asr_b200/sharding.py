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


def gather_records(local: List[Record], rank: int, world: int, group=None, device=None) -> Optional[List[Record]]:
    """The final hypothesis gather: rank 0 receives every shard's records, sorted by utterance index
    (the ``cat | sort -k1`` of ``asr.sh:1603-1612``); other ranks return None.

    The records travel as ONE padded int32 tensor per rank (index, length, fp32 score bits, tokens) through
    ``dist.gather`` (NCCL over NVLink when ``device`` is a CUDA device, gloo on the CPU): one small all-reduce for the
    extents, one gather; no pickling."""
    if world == 1 or not dist.is_initialized():
        return sorted(local, key=lambda r: r[0])
    import numpy as np

    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    ext = torch.tensor([len(local), max((len(r[1]) for r in local), default=0)], dtype=torch.int64, device=device)
    dist.all_reduce(ext, op=dist.ReduceOp.MAX, group=group)
    n_max, l_max = int(ext[0]), int(ext[1])
    pack = np.full((n_max, 3 + l_max), -1, dtype=np.int32)
    for i, (idx, toks, score) in enumerate(local):
        pack[i, 0] = idx
        pack[i, 1] = len(toks)
        pack[i, 2] = np.float32(score).view(np.int32)
        pack[i, 3 : 3 + len(toks)] = toks
    t = torch.from_numpy(pack).to(device)
    out = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
    dist.gather(t, out, dst=0, group=group)
    if rank != 0:
        return None
    merged: List[Record] = []
    for o in out:
        a = o.cpu().numpy()
        for row in a[a[:, 0] >= 0]:
            merged.append((int(row[0]), [int(v) for v in row[3 : 3 + int(row[1])]], float(row[2:3].view(np.float32)[0])))
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

    def collect(b, res):
        if len(res) != len(b):
            raise RuntimeError(f"decode_fn returned {len(res)} results for a batch of {len(b)}")
        for i, (toks, score) in zip(b, res):
            local.append((int(i), [int(t) for t in toks], float(score)))

    # decode_fn may return a callable instead of the results: the batch has been QUEUED on the device and the callable
    # reads it back; it is called after the next batch has been queued, so the read-back of batch i overlaps the decode
    # of batch i + 1
    pending = None
    for b in batches(mine, lengths, batch_size):
        res = decode_fn(b)
        if pending is not None:
            collect(pending[0], pending[1]())
            pending = None
        if callable(res):
            pending = (b, res)
        else:
            collect(b, res)
    if pending is not None:
        collect(pending[0], pending[1]())
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
