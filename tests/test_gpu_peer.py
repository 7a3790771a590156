"""The merge step across ranks over peer memory (peer.cu) against the oracle on the unsplit data.  `-m gpu`.

One process per rank, as in production.  On a box with a single GPU both ranks share device 0 (CUDA IPC works
between processes on one device too; their kernels are time-sliced, so the exchange is slow there, but it is the
same protocol: flags, peer loads, peer stores).  Handles are exchanged over a gloo group on 127.0.0.1.
"""
import os
import socket

import numpy as np
import pytest

import oracle
from tests import util

pytestmark = pytest.mark.gpu

TOL = 1e-11


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _problem(loss, d=257):
    rng = np.random.default_rng(23)
    n = 3000
    row_ptr, col, val = util.random_csr(rng, n, d, 9, skew=1.1, empty_rows=0.03)
    label = util.labels_for(rng, loss, n)
    offset = rng.normal(scale=0.2, size=n)
    weight = rng.uniform(0.5, 2.0, size=n)
    f, s, icpt = util.normalization_context("STANDARDIZATION", n, d, row_ptr, col, val, d - 1)
    w = rng.normal(scale=0.2, size=d)
    v = rng.normal(size=d)
    return n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w, v


def _shards(row_ptr, world, empty_last):
    """Row ranges per rank; with empty_last the last rank gets no rows at all (a straggler partition)."""
    from photon_ml_b200 import parallel
    if not empty_last or world == 1:
        return parallel.shard_rows_by_nnz(row_ptr, world)
    return parallel.shard_rows_by_nnz(row_ptr, world - 1) + [(len(row_ptr) - 1, len(row_ptr) - 1)]


def _worker(rank, world, port, loss, out_dir, multicast, d=257, empty_last=False):
    import torch
    import torch.distributed as dist
    import photon_ml_b200 as pml
    from photon_ml_b200 import parallel

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    device = rank % torch.cuda.device_count()
    torch.cuda.set_device(device)
    n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w, v = _problem(loss, d)
    lo, hi = _shards(row_ptr, world, empty_last)[rank]
    rp, c, vv = parallel.slice_csr(row_ptr, col, val, lo, hi)
    ctx = pml.Context(device)
    ds = pml.Dataset.from_csr(ctx, hi - lo, d, rp, c, vv, label[lo:hi], offset[lo:hi], weight[lo:hi])
    ds.set_normalization(f, s, icpt)
    peer = pml.Peer(ctx, d, multicast=multicast)
    dev = torch.device("cuda", device)
    with torch.cuda.stream(torch.cuda.ExternalStream(ctx.stream, device=dev)):
        coef = torch.from_numpy(w).to(dev)
        vec = torch.from_numpy(v).to(dev)
        value = torch.zeros(1, dtype=torch.float64, device=dev)
        grad = torch.zeros(d, dtype=torch.float64, device=dev)
        hv = torch.zeros(d, dtype=torch.float64, device=dev)
        results = []
        for rep in range(3):   # repeated: the epoch counters advance, the graphs replay
            peer.value_and_gradient(ds, loss, coef.data_ptr(), value.data_ptr(), grad.data_ptr())
            peer.hessian_vector(ds, loss, coef.data_ptr(), vec.data_ptr(), hv.data_ptr(), tag=7)
            ctx.synchronize()
            results.append((float(value.cpu()), grad.cpu().numpy().copy(), hv.cpu().numpy().copy()))
        # the result slab without a copy: [vector, value] of the last call
        peer.value_and_gradient(ds, loss, coef.data_ptr())
        ctx.synchronize()
        slab = torch.as_tensor(peer.result_array(), device=dev).cpu().numpy().copy()
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), value=np.array([r[0] for r in results]),
             grad=np.stack([r[1] for r in results]), hv=np.stack([r[2] for r in results]), slab=slab,
             transport=peer.transport, why=str(getattr(peer, "multicast_error", None)))
    dist.barrier()
    peer.close()
    ds.close()
    ctx.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,multicast", [(1, True), (2, True), (3, True), (2, False)])
def test_peer_merge_matches_the_oracle_on_the_unsplit_data(tmp_path, world, multicast):
    """multicast=True: the NVSwitch multicast transport where the box offers it (ranks on distinct GPUs of one
    NVSwitch fabric), else -- and with multicast=False -- peer-to-peer loads and stores."""
    import torch.multiprocessing as mp
    loss = oracle.LOGISTIC
    mp.spawn(_worker, args=(world, _free_port(), loss, str(tmp_path), multicast), nprocs=world, join=True)
    n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w, v = _problem(loss)
    full = oracle.Data(n, d, row_ptr, col, val, label, offset, weight)
    f_ref, g_ref = oracle.value_and_gradient(full, loss, w, f, s, icpt)
    hv_ref = oracle.hessian_vector(full, loss, w, v, f, s, icpt)
    got = [np.load(str(tmp_path / ("rank%d.npz" % r))) for r in range(world)]
    print("world %d: transport %s (%s)" % (world, got[0]["transport"], got[0]["why"]))
    assert len({str(g["transport"]) for g in got}) == 1
    if not multicast:
        assert str(got[0]["transport"]) == "p2p"
    for r in range(world):
        for rep in range(3):
            assert abs(got[r]["value"][rep] - f_ref) <= TOL * abs(f_ref)
            assert util.rel_err(got[r]["grad"][rep], g_ref) <= TOL
            assert util.rel_err(got[r]["hv"][rep], hv_ref) <= TOL
        assert util.rel_err(got[r]["slab"][:d], g_ref) <= TOL and abs(got[r]["slab"][d] - f_ref) <= TOL * abs(f_ref)
        # one rank sums each element, in rank order, and stores it everywhere: every rank holds the same bits
        assert np.array_equal(got[r]["grad"], got[0]["grad"]) and np.array_equal(got[r]["hv"], got[0]["hv"])
        assert np.array_equal(got[r]["value"], got[0]["value"])
        # ... and repeating the evaluation reproduces them
        assert np.array_equal(got[r]["grad"][0], got[r]["grad"][2])


@pytest.mark.parametrize("d", [2, 5])
def test_peer_merge_tiny_dimension_and_an_empty_shard(tmp_path, d):
    """d + 2 raw sums split over 3 ranks leaves ranks with one pair or none; the last rank holds no rows."""
    import torch.multiprocessing as mp
    loss, world = oracle.POISSON, 3
    mp.spawn(_worker, args=(world, _free_port(), loss, str(tmp_path), False, d, True), nprocs=world, join=True)
    n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w, v = _problem(loss, d)
    full = oracle.Data(n, d, row_ptr, col, val, label, offset, weight)
    f_ref, g_ref = oracle.value_and_gradient(full, loss, w, f, s, icpt)
    hv_ref = oracle.hessian_vector(full, loss, w, v, f, s, icpt)
    for r in range(world):
        got = np.load(str(tmp_path / ("rank%d.npz" % r)))
        assert abs(got["value"][0] - f_ref) <= TOL * abs(f_ref)
        assert util.rel_err(got["grad"][0], g_ref) <= TOL and util.rel_err(got["hv"][2], hv_ref) <= TOL
        assert util.rel_err(got["slab"][:d], g_ref) <= TOL


def test_peer_rejects_mismatched_dimension():
    import photon_ml_b200 as pml
    rng = np.random.default_rng(1)
    n, d = 50, 9
    row_ptr, col, val = util.random_csr(rng, n, d, 3)
    ctx = pml.Context(0)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, util.labels_for(rng, oracle.SQUARED, n))
    peer = pml.Peer(ctx, d + 1, rank=0, world=1)
    with pytest.raises(pml.IllegalArgumentError):
        peer.value_and_gradient(ds, oracle.SQUARED, 1)   # the pointer is never dereferenced
    with pytest.raises(pml.IllegalArgumentError):
        pml.Peer(ctx, d, rank=0, world=9)
    peer.close()
    ds.close()
    ctx.close()
