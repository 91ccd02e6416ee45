"""Row-range sharding of a scan across ranks / GPUs.

Rows are independent, so the table is cut into contiguous ranges, one per rank (one process per GPU); every
rank scans its range with its own query handle and returns the ids of its matching rows; the match lists are
concatenated on the host.  There is NO collective on the data path -- `gather_matches` moves a few hundred
row ids per rank with `torch.distributed.gather_object` (gloo or nccl, whatever the job runs on), and only
because the ranks are separate processes; inside one process (the DuckDB extension, or a C-ABI handle that owns
several devices) the lists are simply appended.

Reference counterpart: DuckDB worker threads mapped round-robin onto devices
(src/cudasp_extension.cu:129-132,516-532) and the mutex-guarded append of matches (:383-391).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np


def shard_range(rows: int, rank: int, world: int, align: int = 1) -> Tuple[int, int]:
    """Contiguous row range [lo, hi) of `rank` out of `world`; boundaries are multiples of `align` (except the end)."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("rank %d out of range for world size %d" % (rank, world))
    per = -(-rows // world)
    if align > 1:
        per = -(-per // align) * align
    lo = min(rows, rank * per)
    hi = min(rows, lo + per)
    return lo, hi


def batch_ranges(lo: int, hi: int, batch_rows: int) -> List[Tuple[int, int]]:
    """The rank's range cut into batches of at most batch_rows rows (the reference's batch_size)."""
    if batch_rows <= 0:
        raise ValueError("batch_rows must be positive")
    return [(b, min(hi, b + batch_rows)) for b in range(lo, hi, batch_rows)]


def local_csr(offsets: np.ndarray, lo: int, hi: int):
    """CSR offsets of rows [lo, hi) rebased to 0, and the [o0, o1) slice of the outputs array they index."""
    o0, o1 = int(offsets[lo]), int(offsets[hi])
    return (offsets[lo: hi + 1].astype(np.int64) - o0).astype(np.uint32), o0, o1


def merge_matches(parts: Sequence[Tuple[int, np.ndarray]]) -> np.ndarray:
    """[(range_lo, local ids)] from every rank / batch -> sorted global row ids (no duplicates expected)."""
    out = [np.asarray(ids, dtype=np.int64) + int(lo) for lo, ids in parts if len(ids)]
    if not out:
        return np.zeros(0, dtype=np.uint32)
    allids = np.sort(np.concatenate(out))
    if len(allids) > 1 and np.any(allids[1:] == allids[:-1]):
        raise ValueError("overlapping shards: a row id was reported twice")
    return allids.astype(np.uint32)


def gather_matches(lo: int, ids: np.ndarray, dist=None, dst: int = 0):
    """Collect every rank's (lo, ids) on `dst` and merge them.  Returns the merged ids on dst, None elsewhere."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return merge_matches([(lo, ids)])
    me, world = dist.get_rank(), dist.get_world_size()
    box = [None] * world if me == dst else None
    dist.gather_object((int(lo), np.asarray(ids, dtype=np.uint32)), box, dst=dst)
    return merge_matches(box) if me == dst else None


def max_over_ranks(seconds: float, dist=None) -> float:
    """Job time = the slowest rank (the bench.py contract)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(seconds)
    import torch
    t = torch.tensor([float(seconds)], dtype=torch.float64)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])
