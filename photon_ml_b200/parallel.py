"""Host-side sharding for one process per GPU (torch.distributed is the plumbing).

Fixed effect: contiguous row shards balanced by entry count; every rank produces the raw sums
[vectorSum(d), valueSum, prefactorSum] of its rows and ONE all-reduce of d+2 doubles replaces
Spark's treeAggregate (lib/function/glm/ValueAndGradientAggregator.scala:243-247 with merge
:161-176); getVector (:199-221) is applied once afterwards.

Random effect: entities are independent (api/algorithm/RandomEffectCoordinate.scala:108-119), so
they are bin-packed over ranks by row count -- the rule of RandomEffectDataSetPartitioner.
fromGameDataSet (api/data/RandomEffectDataSetPartitioner.scala:113-146) -- and no collective runs
on the data path.
"""
import heapq

import numpy as np


def shard_rows_by_nnz(row_ptr, world):
    """Contiguous row ranges [(lo, hi)] * world with (nearly) equal numbers of stored entries."""
    row_ptr = np.asarray(row_ptr, dtype=np.int64)
    n = len(row_ptr) - 1
    nnz = int(row_ptr[-1]) if n > 0 else 0
    cuts = [0]
    for r in range(1, world):
        target = nnz * r // world
        cut = int(np.searchsorted(row_ptr, target, side="left"))
        cuts.append(min(max(cut, cuts[-1]), n))
    cuts.append(n)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def shard_entities(rows_per_entity, world, capacity=None):
    """Greedy bin packing: entities by descending row count, each to the least loaded rank.
    Returns assignment[e] = rank.  Ties resolve to the lowest rank / lowest entity id, so every
    rank computes the same map without communication.  `capacity` is the reference's partitionerCapacity
    (RandomEffectDataSetPartitioner.scala:117, default 10000 there): only that many of the heaviest entities go
    through the heap, the others fall back to the hash partitioner (:74-84; entity id mod world here)."""
    rows_per_entity = np.asarray(rows_per_entity, dtype=np.int64)
    order = np.argsort(-rows_per_entity, kind="stable")
    heap = [(0, r) for r in range(world)]
    heapq.heapify(heap)
    assignment = (np.arange(len(rows_per_entity)) % world).astype(np.int32)
    if capacity is not None:
        order = order[:capacity]
    for e in order:
        load, r = heapq.heappop(heap)
        assignment[e] = r
        heapq.heappush(heap, (load + int(rows_per_entity[e]), r))
    return assignment


def slice_csr(row_ptr, col, val, lo, hi):
    row_ptr = np.asarray(row_ptr, dtype=np.int64)
    a, b = int(row_ptr[lo]), int(row_ptr[hi])
    return row_ptr[lo:hi + 1] - row_ptr[lo], col[a:b], val[a:b]


def all_reduce_sum(tensor):
    """In-place sum over ranks when a process group exists (nccl on GPUs, gloo in the CPU tests)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(tensor, op=dist.ReduceOp.SUM)
    return tensor


def distributed_value_and_gradient(local_raw_fn, finalize_fn, d, device="cpu"):
    """The fixed-effect split used by bench.py and the tests.

    local_raw_fn(out) fills out[0:d] = vectorSum, out[d] = valueSum, out[d+1] = prefactorSum for this
    rank's rows (a torch tensor of d+2 float64 on `device`); finalize_fn(raw) -> (value, vector)
    applies getVector to the reduced sums."""
    import torch
    raw = torch.zeros(d + 2, dtype=torch.float64, device=device)
    local_raw_fn(raw)
    all_reduce_sum(raw)
    return finalize_fn(raw)
