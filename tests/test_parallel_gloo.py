"""world_size-2 CPU (gloo) tests of the host-side multi-GPU logic: row sharding + the raw-sum
all-reduce + single finalisation, and entity bin-packing with no collective.  The per-rank
"kernel" here is the oracle (this is a test of the plumbing, not of the CUDA path)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from photon_ml_b200 import parallel
from tests import util

WORLD = 2


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _problem():
    rng = np.random.default_rng(11)
    n, d = 900, 40
    row_ptr, col, val = util.random_csr(rng, n, d, 6, skew=1.2, empty_rows=0.05)
    label = util.labels_for(rng, 0, n)
    offset = rng.normal(scale=0.3, size=n)
    weight = rng.uniform(0.5, 2.0, size=n)
    f, s, icpt = util.normalization_context("STANDARDIZATION", n, d, row_ptr, col, val, d - 1)
    w = rng.normal(scale=0.3, size=d)
    return n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w


def _raw_sums(data, loss, w, f, s, icpt):
    """[vectorSum, valueSum, prefactorSum] of a shard, recovered from the oracle's finalised output by
    undoing getVector (only used here, to stand in for the device's raw partial sums)."""
    ec = w if f is None else w * f
    shift = 0.0 if s is None else -float(np.dot(ec, s))
    # prefactor = sum wt*l'(z): the gradient w.r.t. a unit-feature column, i.e. gradient of offset
    value, _ = oracle.value_and_gradient(data, loss, w, f, s, icpt)
    # unnormalised sums: evaluate with effective coefficients and shifted offsets, no context
    shifted = oracle.Data(data.n, data.d, data.row_ptr, data.col, data.val, data.label,
                          (data.offset if data.offset is not None else 0.0) + shift + np.zeros(data.n), data.weight)
    value2, vsum = oracle.value_and_gradient(shifted, loss, ec)
    assert abs(value - value2) <= 1e-12 * max(1.0, abs(value))
    ones = oracle.Data(data.n, 1, np.arange(data.n + 1, dtype=np.int64), np.zeros(data.n, np.int32), np.ones(data.n),
                       data.label, shifted.offset + np.array([
                           float(np.dot(data.val[data.row_ptr[i]:data.row_ptr[i + 1]], ec[data.col[data.row_ptr[i]:data.row_ptr[i + 1]]]))
                           for i in range(data.n)]), data.weight)
    _, pref = oracle.value_and_gradient(ones, loss, np.zeros(1))
    return np.concatenate([vsum, [value2, pref[0]]])


def _worker(rank, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w = _problem()
    lo, hi = parallel.shard_rows_by_nnz(row_ptr, WORLD)[rank]
    rp, c, v = parallel.slice_csr(row_ptr, col, val, lo, hi)
    shard = oracle.Data(hi - lo, d, rp, c, v, label[lo:hi], offset[lo:hi], weight[lo:hi])

    def local_raw(out):
        out.copy_(torch.from_numpy(_raw_sums(shard, oracle.LOGISTIC, w, f, s, icpt)))

    def finalize(raw):
        r = raw.numpy()
        return float(r[d]), (r[:d] - s * r[d + 1]) * f      # getVector, ValueAndGradientAggregator.scala:199-221

    value, grad = parallel.distributed_value_and_gradient(local_raw, finalize, d)

    # random effect: entities bin-packed over ranks, no collective on the data path
    rng = np.random.default_rng(3)
    rows_per_entity = rng.integers(1, 200, size=57)
    assign = parallel.shard_entities(rows_per_entity, WORLD)
    mine = np.flatnonzero(assign == rank)
    local_rows = torch.tensor([int(rows_per_entity[mine].sum())], dtype=torch.int64)
    gathered = [torch.zeros(1, dtype=torch.int64) for _ in range(WORLD)]
    dist.all_gather(gathered, local_rows)      # bookkeeping only (results are gathered on the host in GAME)
    if rank == 0:
        np.savez(out_path, value=value, grad=grad, loads=np.array([int(g) for g in gathered]),
                 total=int(rows_per_entity.sum()), n_mine=len(mine))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_row_shard_allreduce_and_entity_packing(tmp_path):
    out_path = str(tmp_path / "rank0.npz")
    mp.spawn(_worker, args=(_free_port(), out_path), nprocs=WORLD, join=True)
    got = np.load(out_path)
    n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w = _problem()
    full = oracle.Data(n, d, row_ptr, col, val, label, offset, weight)
    f_ref, g_ref = oracle.value_and_gradient(full, oracle.LOGISTIC, w, f, s, icpt)
    assert abs(float(got["value"]) - f_ref) <= 1e-12 * abs(f_ref)
    assert util.rel_err(got["grad"], g_ref) <= 1e-12
    loads = got["loads"]
    assert loads.sum() == got["total"]
    assert abs(int(loads[0]) - int(loads[1])) <= 200    # greedy packing: imbalance below the largest entity


def test_shard_rows_by_nnz_balances_entries():
    rng = np.random.default_rng(0)
    lens = rng.integers(0, 50, size=1000)
    lens[100] = 5000
    row_ptr = np.concatenate([[0], np.cumsum(lens)])
    for world in (1, 2, 4, 8):
        shards = parallel.shard_rows_by_nnz(row_ptr, world)
        assert shards[0][0] == 0 and shards[-1][1] == 1000
        assert all(a[1] == b[0] for a, b in zip(shards, shards[1:]))
        sizes = [row_ptr[hi] - row_ptr[lo] for lo, hi in shards]
        assert max(sizes) <= row_ptr[-1] / world + 5000 + 50   # off by at most the longest row


def test_shard_entities_is_deterministic_and_greedy():
    rows = np.array([10, 1, 7, 7, 3, 3, 3, 1])
    a = parallel.shard_entities(rows, 3)
    assert np.array_equal(a, parallel.shard_entities(rows, 3))
    loads = [rows[a == r].sum() for r in range(3)]
    assert sum(loads) == rows.sum() and max(loads) - min(loads) <= 3


def test_shard_entities_capacity_follows_the_reference_partitioner():
    """RandomEffectDataSetPartitioner.fromGameDataSet (api/data/RandomEffectDataSetPartitioner.scala:113-146): only the
    `partitionerCapacity` heaviest entities go through the min-heap; getPartition (:74-84) hashes the others."""
    import numpy as np
    from photon_ml_b200 import parallel
    rng = np.random.default_rng(0)
    rows = np.minimum(rng.zipf(1.8, size=5000), 2000)
    full = parallel.shard_entities(rows, 4)
    capped = parallel.shard_entities(rows, 4, capacity=100)
    heavy = np.argsort(-rows, kind="stable")[:100]
    light = np.setdiff1d(np.arange(5000), heavy)
    assert np.array_equal(capped[light], light % 4)                   # hash fallback
    # the heap part is the same greedy rule as without a cap, applied to the same (heaviest-first) prefix
    assert np.array_equal(capped[heavy], parallel.shard_entities(rows[heavy], 4)[np.arange(100)])
    loads = [rows[full == r].sum() for r in range(4)]
    assert max(loads) - min(loads) <= rows.max()
    loads = [rows[capped == r].sum() for r in range(4)]
    assert max(loads) <= 1.2 * min(loads)
