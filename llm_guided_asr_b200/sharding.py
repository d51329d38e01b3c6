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
    extents, one gather; no pickling.  Same wire format as ``gather_packed`` (a rank may use either)."""
    import numpy as np

    if world == 1 or not dist.is_initialized():
        return sorted(local, key=lambda r: r[0])
    lmax = max((len(r[1]) for r in local), default=0)
    toks = np.full((len(local), lmax), -1, dtype=np.int64)
    for i, r in enumerate(local):
        toks[i, : len(r[1])] = r[1]
    pb = PackedBest(toks, [len(r[1]) for r in local], [r[2] for r in local])
    out = gather_packed([([r[0] for r in local], pb)], rank, world, group, device)
    return None if out is None else out.to_list()


class PackedBest:
    """Best hypotheses of one batch as arrays (what ``FusedStepEngine.gather_best_arrays`` returns): ``tokens (n, L)``
    int, ``lens (n,)``, ``scores (n,)``.  A ``decode_fn`` that returns this (or a callable returning it) keeps the whole
    result path of ``decode_sharded`` vectorised: no per-token Python objects."""

    def __init__(self, tokens, lens, scores):
        import numpy as np

        self.tokens = np.asarray(tokens)
        self.lens = np.asarray(lens, dtype=np.int32)
        self.scores = np.asarray(scores, dtype=np.float32)
        if self.tokens.ndim != 2 or len(self.lens) != len(self.tokens) or len(self.scores) != len(self.tokens):
            raise ValueError("PackedBest: tokens (n, L), lens (n,), scores (n,) expected")

    def __len__(self):
        return len(self.lens)


class PackedRecords:
    """Merged, index-sorted records as arrays; behaves as a sequence of ``(index, tokens, score)`` records."""

    def __init__(self, indices, lens, scores, tokens):
        self.indices, self.lengths, self.scores, self.tokens = indices, lens, scores, tokens

    def __len__(self):
        return len(self.indices)

    def __getitem__(self, i) -> Record:
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        n = int(self.lengths[i])
        return int(self.indices[i]), self.tokens[i, :n].tolist(), float(self.scores[i])

    def __iter__(self):
        return (self[i] for i in range(len(self)))

    def to_list(self) -> List[Record]:
        return list(self)


def _pack_rows(chunks):
    """[(indices, PackedBest)] -> int32 matrix, one row per utterance: index, length, fp32 score bits, tokens (-1 padded)."""
    import numpy as np

    n = sum(len(c[1]) for c in chunks)
    lmax = max((int(c[1].lens.max()) if len(c[1]) else 0 for c in chunks), default=0)
    pack = np.full((n, 3 + lmax), -1, dtype=np.int32)
    r0 = 0
    for idx, pb in chunks:
        m = len(pb)
        w = min(lmax, pb.tokens.shape[1])
        pack[r0 : r0 + m, 0] = np.asarray(idx, dtype=np.int32)
        pack[r0 : r0 + m, 1] = pb.lens
        pack[r0 : r0 + m, 2] = pb.scores.view(np.int32)
        blk = pb.tokens[:, :w].astype(np.int32)
        blk[np.arange(w)[None, :] >= pb.lens[:, None]] = -1
        pack[r0 : r0 + m, 3 : 3 + w] = blk
        r0 += m
    return pack


def gather_packed(chunks, rank: int, world: int, group=None, device=None) -> Optional[PackedRecords]:
    """``gather_records`` on arrays: the same single padded int32 tensor per rank, packed and unpacked with array
    operations only (a 4096-utterance sweep spends tens of ms in per-token Python objects otherwise)."""
    import numpy as np

    pack = _pack_rows(chunks)
    if world > 1 and dist.is_initialized():
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
        ext = torch.tensor([pack.shape[0], pack.shape[1]], dtype=torch.int64, device=device)
        dist.all_reduce(ext, op=dist.ReduceOp.MAX, group=group)
        full = np.full((int(ext[0]), int(ext[1])), -1, dtype=np.int32)
        full[: pack.shape[0], : pack.shape[1]] = pack
        t = torch.from_numpy(full).to(device)
        out = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
        dist.gather(t, out, dst=0, group=group)
        if rank != 0:
            return None
        pack = torch.cat(out).cpu().numpy()
        pack = pack[pack[:, 0] >= 0]
    order = np.argsort(pack[:, 0], kind="stable")
    pack = pack[order]
    return PackedRecords(pack[:, 0].copy(), pack[:, 1].copy(), pack[:, 2].copy().view(np.float32), pack[:, 3:])


def decode_sharded(
    decode_fn: Callable[[List[int]], List[Tuple[List[int], float]]],
    lengths: Sequence[int],
    batch_size: int,
    rank: Optional[int] = None,
    world: Optional[int] = None,
    group=None,
):
    """Decode ``len(lengths)`` utterances over ``world`` ranks.

    ``decode_fn(batch_indices)`` decodes one lock-step batch of this rank's utterances and returns
    ``[(tokens, score)]`` in the order of ``batch_indices`` (on a GPU box:
    ``BatchBeamSearch.decode_batch`` on the encoder outputs of those utterances, best hypothesis) - or a
    ``PackedBest`` (arrays; the vectorised path), or a callable returning either (deferred read-back, below).
    Returns the merged, index-sorted records on rank 0 (a list, or ``PackedRecords`` on the array path), None
    elsewhere."""
    if rank is None:
        rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world is None:
        world = dist.get_world_size(group) if dist.is_initialized() else 1
    mine = partition(lengths, world)[rank]
    local: List[Record] = []
    chunks = []

    def collect(b, res):
        if len(res) != len(b):
            raise RuntimeError(f"decode_fn returned {len(res)} results for a batch of {len(b)}")
        if isinstance(res, PackedBest):
            chunks.append((list(b), res))
            return
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
    if chunks and local:
        raise RuntimeError("decode_fn mixed PackedBest and list results")
    if chunks:
        return gather_packed(chunks, rank, world, group)
    # (a rank of a packed run whose shard is empty lands here with nothing: it still has to take part in the gather)
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
