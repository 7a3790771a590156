"""The merge step across ranks over peer memory (peer.cu) against the oracle on the unsplit data.  `-m gpu`.

Two ways in, the same kernels behind both:
  * one process per rank, as in production, when the box has at least as many GPUs as ranks (handles exchanged over
    a gloo group on 127.0.0.1);
  * pml_peer_emulate for every world size up to 8 on ONE GPU: all ranks' CTAs run inside one cooperative launch, so
    the flag waits between ranks cannot starve.  (Ranks as separate processes time-slicing one GPU are not run:
    kernels that wait on one another must not share a GPU as separate launches.)
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


def _need_gpus(world):
    import torch
    if world > 1 and torch.cuda.device_count() < world:
        pytest.skip("%d ranks need %d GPUs (one process per GPU); the exchange kernels are covered on one GPU by the "
                    "pml_peer_emulate tests below" % (world, world))


@pytest.mark.parametrize("world,multicast", [(1, True), (2, True), (3, True), (2, False), (4, False), (8, False)])
def test_peer_merge_matches_the_oracle_on_the_unsplit_data(tmp_path, world, multicast):
    """multicast=True: the NVSwitch multicast transport where the box offers it (ranks on distinct GPUs of one
    NVSwitch fabric), else -- and with multicast=False -- peer-to-peer loads and stores."""
    import torch.multiprocessing as mp
    _need_gpus(world)
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
    _need_gpus(world)
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


def _host_worker(rank, world, port, loss, out_dir):
    """The host-buffer entry points: every rank uploads its slice, the GPUs exchange slices, every rank downloads its
    slice of the result into ONE shared buffer (a memory-mapped file)."""
    import torch
    import torch.distributed as dist
    import photon_ml_b200 as pml
    from photon_ml_b200 import parallel

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    device = rank % torch.cuda.device_count()
    torch.cuda.set_device(device)
    n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w, v = _problem(loss)
    lo, hi = _shards(row_ptr, world, False)[rank]
    rp, c, vv = parallel.slice_csr(row_ptr, col, val, lo, hi)
    ctx = pml.Context(device)
    ds = pml.Dataset.from_csr(ctx, hi - lo, d, rp, c, vv, label[lo:hi], offset[lo:hi], weight[lo:hi])
    ds.set_normalization(f, s, icpt)
    peer = pml.Peer(ctx, d, multicast=False)
    shared_g = np.memmap(os.path.join(out_dir, "grad.bin"), dtype=np.float64, mode="r+", shape=(d,))
    shared_h = np.memmap(os.path.join(out_dir, "hv.bin"), dtype=np.float64, mode="r+", shape=(d,))
    full_g, full_h = np.zeros(d), np.zeros(d)
    values = []
    for rep in range(3):
        values.append(peer.value_and_gradient_host(ds, loss, w, shared_g))
        peer.hessian_vector_host(ds, loss, w, v, shared_h, tag=9)
        peer.hessian_vector_host(ds, loss, None, v, shared_h, tag=9)      # cached pass: no coefficients needed
    values.append(peer.value_and_gradient_host(ds, loss, w, full_g, gather_output=True))
    peer.hessian_vector_host(ds, loss, w, v, full_h, gather_output=True)
    full_d = np.zeros(d)
    peer.hessian_diagonal_host(ds, loss, w, full_d, gather_output=True)
    shared_g.flush()
    shared_h.flush()
    np.savez(os.path.join(out_dir, "host%d.npz" % rank), values=np.array(values), grad=full_g, hv=full_h, diag=full_d,
             slice=np.array(peer.slice()), status=peer.status())
    dist.barrier()
    peer.close()
    ds.close()
    ctx.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_host_buffer_allreduce_entry_points(tmp_path, world):
    import torch.multiprocessing as mp
    _need_gpus(world)
    loss = oracle.LOGISTIC
    n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w, v = _problem(loss)
    for name in ("grad.bin", "hv.bin"):
        np.zeros(d).tofile(str(tmp_path / name))
    mp.spawn(_host_worker, args=(world, _free_port(), loss, str(tmp_path)), nprocs=world, join=True)
    full = oracle.Data(n, d, row_ptr, col, val, label, offset, weight)
    f_ref, g_ref = oracle.value_and_gradient(full, loss, w, f, s, icpt)
    hv_ref = oracle.hessian_vector(full, loss, w, v, f, s, icpt)
    diag_ref = oracle.hessian_diagonal(full, loss, w)     # (no normalization in the diagonal, as in the reference)
    shared_g, shared_h = np.fromfile(str(tmp_path / "grad.bin")), np.fromfile(str(tmp_path / "hv.bin"))
    assert util.rel_err(shared_g, g_ref) <= TOL and util.rel_err(shared_h, hv_ref) <= TOL
    covered = np.zeros(d, dtype=int)
    for r in range(world):
        got = np.load(str(tmp_path / ("host%d.npz" % r)))
        assert got["status"] == 0
        assert util.rel_err(got["diag"], diag_ref) <= TOL
        assert np.all(np.abs(got["values"] - f_ref) <= TOL * abs(f_ref))
        assert np.array_equal(got["grad"], shared_g)           # same path, same bits on every rank
        assert util.rel_err(got["hv"], shared_h) <= 1e-13      # margin pass recomputed vs cached curvature
        covered[got["slice"][0]:got["slice"][1]] += 1
    assert np.all(covered == 1)   # the ranks' slices tile [0, d)


def _shard_raw_sums(ctx, loss, world, d=257, empty_last=False):
    """Every rank's raw sums [vectorSum(d), valueSum, prefactorSum] of value+gradient and of a Hessian-vector product,
    evaluated shard by shard on this one GPU."""
    import torch
    import photon_ml_b200 as pml
    from photon_ml_b200 import parallel
    n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w, v = _problem(loss, d)
    dev = torch.device("cuda", ctx.device)
    vg, hv = np.zeros((world, d + 2)), np.zeros((world, d + 2))
    with torch.cuda.stream(torch.cuda.ExternalStream(ctx.stream, device=dev)):
        coef, vec = torch.from_numpy(w).to(dev), torch.from_numpy(v).to(dev)
        raw = torch.zeros(d + 2, dtype=torch.float64, device=dev)
        for rank, (lo, hi) in enumerate(_shards(row_ptr, world, empty_last)):
            rp, c, vv = parallel.slice_csr(row_ptr, col, val, lo, hi)
            ds = pml.Dataset.from_csr(ctx, hi - lo, d, rp, c, vv, label[lo:hi], offset[lo:hi], weight[lo:hi])
            ds.set_normalization(f, s, icpt)
            ds.value_and_gradient_raw_dev(loss, coef.data_ptr(), raw.data_ptr())
            ctx.synchronize()
            vg[rank] = raw.cpu().numpy()
            ds.hessian_vector_raw_dev(loss, coef.data_ptr(), vec.data_ptr(), raw.data_ptr())
            ctx.synchronize()
            hv[rank] = raw.cpu().numpy()
            ds.close()
    return vg, hv


@pytest.fixture(scope="module")
def ctx():
    import photon_ml_b200 as pml
    c = pml.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("world", [1, 2, 3, 4, 5, 8])
def test_emulated_exchange_matches_the_oracle_on_the_unsplit_data(ctx, world):
    """peer_merge_kernel<world> and peer_gather_kernel<world>, all ranks in one cooperative launch on this GPU."""
    from photon_ml_b200 import abi
    loss = oracle.LOGISTIC
    n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w, v = _problem(loss)
    vg, hv = _shard_raw_sums(ctx, loss, world)
    full = oracle.Data(n, d, row_ptr, col, val, label, offset, weight)
    f_ref, g_ref = oracle.value_and_gradient(full, loss, w, f, s, icpt)
    hv_ref = oracle.hessian_vector(full, loss, w, v, f, s, icpt)
    rng = np.random.default_rng(world)
    vec_in = rng.normal(size=(world, d))        # rank q only contributes ITS slice of row q
    res, gathered = abi.peer_emulate(ctx, vg, f, s, vec_in=vec_in, repeats=3)
    res_hv = abi.peer_emulate(ctx, hv, f, s)
    for r in range(world):
        assert abs(res[r, d] - f_ref) <= TOL * abs(f_ref)
        assert util.rel_err(res[r, :d], g_ref) <= TOL
        assert util.rel_err(res_hv[r, :d], hv_ref) <= TOL
        assert np.array_equal(res[r], res[0]) and np.array_equal(res_hv[r], res_hv[0])   # same bits on every rank
    # the all-gather: every rank ends up with the concatenation of the ranks' slices
    per = ((d + 1) // 2 + world - 1) // world
    want = np.concatenate([vec_in[q, min(2 * per * q, d):min(2 * per * (q + 1), d)] for q in range(world)])
    for r in range(world):
        assert np.array_equal(gathered[r], want)


@pytest.mark.parametrize("world,d", [(3, 2), (3, 5), (8, 2), (8, 13), (4, 1_000_003), (8, 1_000_003)])
def test_emulated_exchange_sums_in_rank_order_at_any_dimension(ctx, world, d):
    """Tiny vectors (ranks left with one pair or none) and the bench dimension: the result is the ranks' raw sums
    added in rank order, then getVector -- reproduced bit for bit with numpy."""
    from photon_ml_b200 import abi
    rng = np.random.default_rng(1000 * world + d % 1000)
    raw = rng.normal(size=(world, d + 2))
    factors = rng.uniform(0.5, 2.0, size=d)
    shifts = rng.normal(size=d)
    vec_in = rng.normal(size=(world, d))
    res, gathered = abi.peer_emulate(ctx, raw, factors, shifts, vec_in=vec_in, repeats=2)
    total = np.zeros(d + 2)
    for q in range(world):
        total = total + raw[q]
    want = np.empty(d + 1)
    want[:d] = (total[:d] - shifts * total[d + 1]) * factors
    want[d] = total[d]
    for r in range(world):
        assert np.array_equal(res[r], want)
    per = ((d + 1) // 2 + world - 1) // world
    cat = np.concatenate([vec_in[q, min(2 * per * q, d):min(2 * per * (q + 1), d)] for q in range(world)])
    for r in range(world):
        assert np.array_equal(gathered[r], cat)
    # without a NormalizationContext
    res = abi.peer_emulate(ctx, raw)
    assert np.array_equal(res[world - 1, :d], total[:d]) and res[0, d] == total[d]


def test_emulated_exchange_with_an_empty_shard(ctx):
    from photon_ml_b200 import abi
    loss, world, d = oracle.POISSON, 4, 5
    n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w, v = _problem(loss, d)
    vg, _ = _shard_raw_sums(ctx, loss, world, d, empty_last=True)
    assert np.all(vg[world - 1] == 0.0)
    full = oracle.Data(n, d, row_ptr, col, val, label, offset, weight)
    f_ref, g_ref = oracle.value_and_gradient(full, loss, w, f, s, icpt)
    res = abi.peer_emulate(ctx, vg, f, s)
    assert abs(res[2, d] - f_ref) <= TOL * abs(f_ref) and util.rel_err(res[2, :d], g_ref) <= TOL


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


def _fe_worker(rank, world, port, out_dir):
    """A fixed-effect coordinate trained SPMD: every rank holds a row shard and runs the same TRON on merged numbers."""
    import torch
    import torch.distributed as dist
    import photon_ml_b200 as pml
    from photon_ml_b200 import parallel
    from photon_ml_b200.game import FixedEffect

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    device = rank % torch.cuda.device_count()
    torch.cuda.set_device(device)
    n, d, row_ptr, col, val, label, offset, weight, f, s, icpt, w, v = _problem(oracle.LOGISTIC)
    lo, hi = _shards(row_ptr, world, False)[rank]
    rp, c, vv = parallel.slice_csr(row_ptr, col, val, lo, hi)
    ctx = pml.Context(device)
    ds = pml.Dataset.from_csr(ctx, hi - lo, d, rp, c, vv, label[lo:hi], offset[lo:hi], weight[lo:hi])
    peer = pml.Peer(ctx, d, multicast=False) if world > 1 else None
    fe = FixedEffect(ds, pml.LOGISTIC, l2RegWeight=0.7, peer=peer)
    with fe.objective.stream():
        value, info = fe.train()
        model = fe.model.cpu().numpy()
        # another coordinate's scores arrive as new offsets (Coordinate.updateModel): re-train from the current model
        fe.set_offsets(torch.from_numpy(offset[lo:hi] * -1.5 + 0.3).to(torch.device("cuda", device)))
        value2, info2 = fe.train()
        model2 = fe.model.cpu().numpy()
    np.savez(os.path.join(out_dir, "fe%d_of_%d.npz" % (rank, world)), model=model, value=value,
             iterations=info["iterations"], hv=info["hessian_vector_evaluations"], model2=model2, value2=value2,
             iterations2=info2["iterations"])
    dist.barrier()
    if peer is not None:
        peer.close()
    ds.close()
    ctx.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_fixed_effect_trains_the_same_model_as_one_gpu(tmp_path, world):
    import torch.multiprocessing as mp
    _need_gpus(world)
    mp.spawn(_fe_worker, args=(1, _free_port(), str(tmp_path)), nprocs=1, join=True)
    mp.spawn(_fe_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    one = np.load(str(tmp_path / "fe0_of_1.npz"))
    for r in range(world):
        got = np.load(str(tmp_path / ("fe%d_of_%d.npz" % (r, world))))
        assert got["iterations"] == one["iterations"] and got["hv"] == one["hv"]
        assert util.rel_err(got["model"], one["model"]) <= 1e-9
        assert abs(got["value"] - one["value"]) <= 1e-11 * abs(one["value"])
        assert got["iterations2"] == one["iterations2"] and util.rel_err(got["model2"], one["model2"]) <= 1e-9
        assert abs(got["value2"] - one["value2"]) <= 1e-11 * abs(one["value2"])
        # every rank trained on the same merged numbers: the same model to the last bit
        assert np.array_equal(got["model"], np.load(str(tmp_path / ("fe0_of_%d.npz" % world)))["model"])
