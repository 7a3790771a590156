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
