"""Coordinate descent with device-resident scores / offsets (photon_ml_b200/game.py) against the same passes done on
the host with the CPU oracle: fixed effect + per-user + per-item random effects, the configs[3] (GLMix) shape in
miniature.  `-m gpu`."""
import numpy as np
import pytest

import oracle
from tests import util
from tests.objectives import OracleGLMObjective

pytestmark = pytest.mark.gpu


def _grouped(rng, n, entity_of_row, d_re, k):
    """Rows grouped by entity: (row_map, ent_row_ptr, csr in grouped order)."""
    n_ent = int(entity_of_row.max()) + 1
    row_map = np.argsort(entity_of_row, kind="stable").astype(np.int64)
    rows = np.bincount(entity_of_row, minlength=n_ent)
    ent_row_ptr = np.concatenate([[0], np.cumsum(rows)]).astype(np.int64)
    ptr, col, val = util.random_csr(rng, n, d_re, k)
    return row_map, ent_row_ptr, ptr, col, val


def _host_reference(loss, l2, n, label, base, fe, res, passes):
    """The same coordinate descent with numpy vectors and the oracle objective, entity by entity."""
    from photon_ml_b200.function import NoNormalization
    from photon_ml_b200.optimization import TRON
    d_fe, fe_ptr, fe_col, fe_val = fe
    w_fe = np.zeros(d_fe)
    w_re = [np.zeros(int(len(r["ent_row_ptr"]) - 1) * r["d"]) for r in res]
    scores = [np.zeros(n) for _ in range(1 + len(res))]
    history = []
    for _ in range(passes):
        partial = sum(scores[1:])
        data = oracle.Data(n, d_fe, fe_ptr, fe_col, fe_val, label, base + partial)
        opt = TRON(NoNormalization())
        w_fe, f = opt.optimize(OracleGLMObjective(loss, l2RegWeight=l2), w_fe, data)
        scores[0] = util.csr_to_dense(n, d_fe, fe_ptr, fe_col, fe_val) @ w_fe
        history.append(f + sum(0.5 * l2 * w.dot(w) for w in w_re))
        for c, r in enumerate(res):
            partial = sum(s for i, s in enumerate(scores) if i != c + 1)
            off = (base + partial)[r["row_map"]]
            lab = label[r["row_map"]]
            total, local_scores = 0.0, np.zeros(n)
            for e in range(len(r["ent_row_ptr"]) - 1):
                lo, hi = int(r["ent_row_ptr"][e]), int(r["ent_row_ptr"][e + 1])
                sl = slice(int(r["ptr"][lo]), int(r["ptr"][hi]))
                local = oracle.Data(hi - lo, r["d"], r["ptr"][lo:hi + 1] - r["ptr"][lo], r["col"][sl], r["val"][sl],
                                    lab[lo:hi], off[lo:hi])
                opt = TRON(NoNormalization())
                seg = slice(e * r["d"], (e + 1) * r["d"])
                w_re[c][seg], f = opt.optimize(OracleGLMObjective(loss, l2RegWeight=l2), w_re[c][seg], local)
                total += f
                X = util.csr_to_dense(hi - lo, r["d"], local.row_ptr, local.col, local.val)
                local_scores[lo:hi] = X @ w_re[c][seg]
            scores[c + 1] = np.zeros(n)
            scores[c + 1][r["row_map"]] = local_scores
            history.append(total + 0.5 * l2 * w_fe.dot(w_fe) + sum(0.5 * l2 * w.dot(w) for i, w in enumerate(w_re) if i != c))
    return w_fe, w_re, history


def test_glmix_shaped_coordinate_descent_matches_the_host_passes():
    import torch
    import photon_ml_b200 as pml
    from photon_ml_b200.game import CoordinateDescent, FixedEffect, RandomEffect
    loss, l2, passes = oracle.LOGISTIC, 1.0, 2
    rng = np.random.default_rng(12)
    n, d_fe = 3000, 40
    fe_ptr, fe_col, fe_val = util.random_csr(rng, n, d_fe, 8)
    label = (rng.random(n) < 0.4).astype(np.float64)
    base = rng.normal(scale=0.1, size=n)
    users, items = rng.integers(0, 120, size=n), rng.integers(0, 50, size=n)
    res = []
    for ent, d_re, k in ((users, 6, 3), (items, 5, 3)):
        row_map, ent_row_ptr, ptr, col, val = _grouped(rng, n, ent, d_re, k)
        res.append({"row_map": row_map, "ent_row_ptr": ent_row_ptr, "ptr": ptr, "col": col, "val": val, "d": d_re})
    ctx = pml.Context(0)
    dev = torch.device("cuda", 0)
    fe_ds = pml.Dataset.from_csr(ctx, n, d_fe, fe_ptr, fe_col, fe_val, label)
    coords = {"fixed": FixedEffect(fe_ds, loss, l2RegWeight=l2, base_offset=torch.tensor(base, device=dev))}
    batches = []
    for name, r in zip(("per-user", "per-item"), res):
        n_ent = len(r["ent_row_ptr"]) - 1
        b = pml.Batch(ctx, r["ent_row_ptr"], np.full(n_ent, r["d"], np.int32), r["ptr"], r["col"], r["val"], label[r["row_map"]])
        batches.append(b)
        coords[name] = RandomEffect(b, torch.tensor(r["row_map"], device=dev), loss, l2RegWeight=l2,
                                    base_offset=torch.tensor(base[r["row_map"]], device=dev))
    models, history = CoordinateDescent(coords, n, descent_iterations=passes).run()
    w_fe, w_re, ref_history = _host_reference(loss, l2, n, label, base, (d_fe, fe_ptr, fe_col, fe_val), res, passes)
    assert len(history) == len(ref_history) == passes * 3
    for h, f_ref in zip(history, ref_history):
        assert abs(h["objective"] - f_ref) <= 1e-8 * abs(f_ref), h
    assert [h["objective"] for h in history] == sorted([h["objective"] for h in history], reverse=True)   # descent
    assert np.max(np.abs(models["fixed"].cpu().numpy() - w_fe)) <= 1e-6
    for name, w in zip(("per-user", "per-item"), w_re):
        assert np.max(np.abs(models[name].cpu().numpy() - w)) <= 1e-6
    for b in batches:
        b.close()
    fe_ds.close()
    ctx.close()


def _sharded_problem():
    """Fixed effect + per-user + per-item effects on 2400 rows, global row order = user-major."""
    rng = np.random.default_rng(77)
    n, n_users, n_items, d_fe = 2400, 60, 25, 40
    user_of_row = np.sort(rng.integers(0, n_users, size=n))            # rows grouped by user already
    item_of_row = rng.integers(0, n_items, size=n)
    fe_ptr, fe_col, fe_val = util.random_csr(rng, n, d_fe, 6, skew=1.1)
    u_ptr, u_col, u_val = util.random_csr(rng, n, 5, 3)               # per-user features, global (user-major) order
    i_ptr, i_col, i_val = util.random_csr(rng, n, 4, 2)               # per-item features, global order
    label = util.labels_for(rng, oracle.LOGISTIC, n)
    return n, n_users, n_items, d_fe, user_of_row, item_of_row, (fe_ptr, fe_col, fe_val), (u_ptr, u_col, u_val), \
        (i_ptr, i_col, i_val), label


def _take_rows(csr, rows):
    """CSR restricted to `rows` (in that order)."""
    ptr, col, val = csr
    lens = (ptr[1:] - ptr[:-1])[rows]
    new_ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    idx = np.concatenate([np.arange(ptr[r], ptr[r + 1]) for r in rows]) if len(rows) else np.zeros(0, dtype=np.int64)
    return new_ptr, col[idx.astype(np.int64)], val[idx.astype(np.int64)]


def _cd_worker(rank, world, port, out_dir):
    import os
    import torch
    import torch.distributed as dist
    import photon_ml_b200 as pml
    from photon_ml_b200 import parallel
    from photon_ml_b200.game import CoordinateDescent, FixedEffect, RandomEffect

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    device = rank % torch.cuda.device_count()
    torch.cuda.set_device(device)
    dev = torch.device("cuda", device)
    n, n_users, n_items, d_fe, user_of_row, item_of_row, fe, ufeat, ifeat, label = _sharded_problem()
    ctx = pml.Context(device)
    # users -> ranks, a rank owns its users' rows (fixed effect + per-user effect); items -> ranks separately
    my_users = np.flatnonzero(parallel.shard_entities(np.bincount(user_of_row, minlength=n_users), world) == rank)
    my_rows = np.flatnonzero(np.isin(user_of_row, my_users)).astype(np.int64)
    my_items = np.flatnonzero(parallel.shard_entities(np.bincount(item_of_row, minlength=n_items), world) == rank)
    item_rows = np.flatnonzero(np.isin(item_of_row, my_items)).astype(np.int64)
    item_rows = item_rows[np.argsort(item_of_row[item_rows], kind="stable")]          # grouped by item
    fp, fc, fv = _take_rows(fe, my_rows)
    fe_ds = pml.Dataset.from_csr(ctx, len(my_rows), d_fe, fp, fc, fv, label[my_rows])
    up, uc, uv = _take_rows(ufeat, my_rows)
    u_ptr = np.concatenate([[0], np.cumsum(np.bincount(user_of_row[my_rows], minlength=n_users)[my_users])]).astype(np.int64)
    user_batch = pml.Batch(ctx, u_ptr, np.full(len(my_users), 5, np.int32), up, uc, uv, label[my_rows])
    ip, ic, iv = _take_rows(ifeat, item_rows)
    i_ptr = np.concatenate([[0], np.cumsum(np.bincount(item_of_row[item_rows], minlength=n_items)[my_items])]).astype(np.int64)
    item_batch = pml.Batch(ctx, i_ptr, np.full(len(my_items), 4, np.int32), ip, ic, iv, label[item_rows])
    peer = pml.Peer(ctx, d_fe, multicast=False) if world > 1 else None

    def all_sum(x):
        t = torch.tensor([x], dtype=torch.float64)
        dist.all_reduce(t)
        return float(t)

    def exchange(t):
        h = t.cpu()
        dist.all_reduce(h)
        t.copy_(h)

    rows_t, items_t = torch.from_numpy(my_rows).to(dev), torch.from_numpy(item_rows).to(dev)
    coords = {"fixed": FixedEffect(fe_ds, pml.LOGISTIC, l2RegWeight=0.8, row_map=rows_t, peer=peer),
              "per-user": RandomEffect(user_batch, rows_t, pml.LOGISTIC, l2RegWeight=0.8),
              "per-item": RandomEffect(item_batch, items_t, pml.LOGISTIC, l2RegWeight=0.8)}
    cd = CoordinateDescent(coords, n, descent_iterations=2, all_reduce_sum=all_sum if world > 1 else None,
                           exchange_scores=exchange if world > 1 else None)
    models, history = cd.run()
    # per-user models keyed by user id, so that runs with different shardings can be compared
    user_models = {int(u): models["per-user"][5 * k:5 * k + 5].cpu().numpy() for k, u in enumerate(my_users)}
    np.savez(os.path.join(out_dir, "cd%d_of_%d.npz" % (rank, world)), fixed=models["fixed"].cpu().numpy(),
             objective=np.array([h["objective"] for h in history]), users=np.array(sorted(user_models)),
             user_models=np.stack([user_models[u] for u in sorted(user_models)]) if user_models else np.zeros((0, 5)))
    dist.barrier()
    if peer is not None:
        peer.close()
    ctx.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_coordinate_descent_matches_one_gpu(tmp_path, world):
    """One process per GPU: users and items bin-packed over the ranks, the fixed effect merged through the peer
    exchange, score vectors summed over the ranks after every update -- the same objective history and models as the
    whole problem on one GPU."""
    import socket
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < world:
        pytest.skip("%d ranks need %d GPUs (one process per GPU)" % (world, world))

    def port():
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        p = s.getsockname()[1]
        s.close()
        return p

    mp.spawn(_cd_worker, args=(1, port(), str(tmp_path)), nprocs=1, join=True)
    mp.spawn(_cd_worker, args=(world, port(), str(tmp_path)), nprocs=world, join=True)
    one = np.load(str(tmp_path / "cd0_of_1.npz"))
    seen = {}
    for r in range(world):
        got = np.load(str(tmp_path / ("cd%d_of_%d.npz" % (r, world))))
        assert np.max(np.abs(got["objective"] - one["objective"]) / np.abs(one["objective"])) <= 1e-9
        assert util.rel_err(got["fixed"], one["fixed"]) <= 1e-7
        for u, m in zip(got["users"], got["user_models"]):
            seen[int(u)] = m
    assert sorted(seen) == list(one["users"])
    for u, m in zip(one["users"], one["user_models"]):
        assert np.max(np.abs(seen[int(u)] - m)) <= 1e-6 * max(1.0, np.max(np.abs(m)))
