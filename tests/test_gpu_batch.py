"""Batched per-entity (random-effect) evaluation against the oracle run entity by entity --
what RandomEffectCoordinate.updateModel does with one SingleNodeGLMLossFunction per entity
(api/algorithm/RandomEffectCoordinate.scala:108-119).  `-m gpu`."""
import numpy as np
import pytest

import oracle
from tests import util

pytestmark = pytest.mark.gpu
TOL = 1e-11


def _entities(rng, n_entities, max_dim, mean_rows, skew_rows=True):
    """Concatenated per-entity CSR in local column spaces; power-law row counts, some empty entities."""
    if skew_rows:
        rows = np.minimum((rng.pareto(1.2, size=n_entities) * mean_rows / 3 + 1).astype(np.int64), 40 * mean_rows)
    else:
        rows = rng.integers(1, 2 * mean_rows, size=n_entities)
    rows[rng.random(n_entities) < 0.03] = 0          # entities with no active rows
    dims = rng.integers(1, max_dim + 1, size=n_entities).astype(np.int32)
    ent_row_ptr = np.concatenate([[0], np.cumsum(rows)]).astype(np.int64)
    row_ptr, col, val = [0], [], []
    for e in range(n_entities):
        d_e = int(dims[e])
        for _ in range(int(rows[e])):
            k = int(rng.integers(0, min(d_e, 8) + 1))   # rows may be empty too
            c = np.sort(rng.choice(d_e, size=k, replace=False)) if k else np.zeros(0, dtype=np.int64)
            col.extend(c.tolist())
            val.extend((rng.integers(1, 9, size=k) / 4.0).tolist())
            row_ptr.append(len(col))
    n = int(ent_row_ptr[-1])
    return (ent_row_ptr, dims, np.asarray(row_ptr, np.int64), np.asarray(col, np.int32), np.asarray(val, np.float64), n)


@pytest.mark.parametrize("loss", [oracle.LOGISTIC, oracle.POISSON, oracle.SQUARED])
def test_batch_matches_per_entity_oracle(loss):
    import photon_ml_b200 as pml
    rng = np.random.default_rng(5 + loss)
    ent_row_ptr, dims, row_ptr, col, val, n = _entities(rng, 600, 24, 12)
    label = util.labels_for(rng, loss, n)
    offset = rng.normal(scale=0.2, size=n)
    weight = rng.uniform(0.5, 2.0, size=n)
    ctx = pml.Context(0)
    batch = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, col, val, label, offset, weight)
    assert batch.d_total == int(dims.sum())
    coefs = rng.normal(scale=0.3, size=batch.d_total)
    vs = rng.normal(size=batch.d_total)
    values, grads = batch.value_and_gradient(loss, coefs)
    hvs = batch.hessian_vector(loss, coefs, vs)
    for e in range(len(dims)):
        lo, hi = int(ent_row_ptr[e]), int(ent_row_ptr[e + 1])
        a, b = int(batch.coef_ptr[e]), int(batch.coef_ptr[e + 1])
        sl = slice(int(row_ptr[lo]), int(row_ptr[hi]))
        local = oracle.Data(hi - lo, int(dims[e]), row_ptr[lo:hi + 1] - row_ptr[lo], col[sl], val[sl], label[lo:hi],
                            offset[lo:hi], weight[lo:hi])
        f_ref, g_ref = oracle.value_and_gradient(local, loss, coefs[a:b])
        assert abs(values[e] - f_ref) <= TOL * max(abs(f_ref), 1e-300), (e, values[e], f_ref)
        assert util.rel_err(grads[a:b], g_ref) <= TOL, e
        assert util.rel_err(hvs[a:b], oracle.hessian_vector(local, loss, coefs[a:b], vs[a:b])) <= TOL, e
    # residual offsets are refreshed through the underlying data set (Coordinate.scala:59-62)
    off2 = rng.normal(size=n)
    batch.dataset.set_offsets(off2)
    values2, _ = batch.value_and_gradient(loss, coefs, want_gradient=False)
    e = int(np.argmax(np.diff(ent_row_ptr)))
    lo, hi = int(ent_row_ptr[e]), int(ent_row_ptr[e + 1])
    sl = slice(int(row_ptr[lo]), int(row_ptr[hi]))
    local = oracle.Data(hi - lo, int(dims[e]), row_ptr[lo:hi + 1] - row_ptr[lo], col[sl], val[sl], label[lo:hi],
                        off2[lo:hi], weight[lo:hi])
    a, b = int(batch.coef_ptr[e]), int(batch.coef_ptr[e + 1])
    f_ref, _ = oracle.value_and_gradient(local, loss, coefs[a:b])
    assert abs(values2[e] - f_ref) <= TOL * abs(f_ref)
    batch.close()
    ctx.close()


def test_batch_rejects_bad_local_columns():
    import photon_ml_b200 as pml
    ctx = pml.Context(0)
    ent_row_ptr = np.array([0, 2, 3], np.int64)
    dims = np.array([2, 3], np.int32)
    row_ptr = np.array([0, 1, 2, 3], np.int64)
    with pytest.raises(pml.IllegalArgumentError):   # column 2 does not exist in entity 0 (d_e = 2)
        pml.Batch(ctx, ent_row_ptr, dims, row_ptr, np.array([0, 2, 1], np.int32), np.ones(3), np.zeros(3))
    b = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, np.array([0, 1, 2], np.int32), np.ones(3), np.zeros(3))
    values, grads = b.value_and_gradient(oracle.LOGISTIC, np.zeros(5))
    assert np.allclose(values, [2 * np.log(2.0), np.log(2.0)], rtol=1e-14)
    assert np.allclose(grads, [0.5, 0.5, 0.0, 0.0, 0.5], rtol=1e-14)
    b.close()
    ctx.close()


def test_batch_device_buffers_match_host_buffers():
    import torch
    import photon_ml_b200 as pml
    rng = np.random.default_rng(77)
    ent_row_ptr, dims, row_ptr, col, val, n = _entities(rng, 900, 16, 20)
    label = util.labels_for(rng, 0, n)
    ctx = pml.Context(0)
    batch = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, col, val, label)
    coefs = rng.normal(scale=0.3, size=batch.d_total)
    values, grads = batch.value_and_gradient(pml.LOGISTIC, coefs)
    dev = torch.device("cuda", 0)
    t_coef = torch.tensor(coefs, device=dev)
    t_val = torch.full((len(dims),), float("nan"), dtype=torch.float64, device=dev)
    t_grad = torch.full((batch.d_total,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    batch.value_and_gradient_dev(pml.LOGISTIC, t_coef.data_ptr(), t_val.data_ptr(), t_grad.data_ptr())
    ctx.synchronize()
    assert np.array_equal(t_val.cpu().numpy(), values)       # same kernels, same order: bit-identical
    assert np.array_equal(t_grad.cpu().numpy(), grads)
    batch.close()
    ctx.close()


def test_batch_active_mask_and_batch_offsets():
    """pml_batch_set_active takes entities out of value / gradient / Hessian-vector results (host and device
    buffers); pml_batch_set_offsets is pml_dataset_set_offsets on the concatenated rows."""
    import torch
    import photon_ml_b200 as pml
    loss = oracle.LOGISTIC
    rng = np.random.default_rng(77)
    ent_row_ptr, dims, row_ptr, col, val, n = _entities(rng, 300, 16, 10)
    label = util.labels_for(rng, loss, n)
    ctx = pml.Context(0)
    batch = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, col, val, label)
    coefs = rng.normal(scale=0.3, size=batch.d_total)
    vs = rng.normal(size=batch.d_total)
    values, grads = batch.value_and_gradient(loss, coefs)
    batch.hessian_vector(loss, coefs, vs, tag=3)
    hvs = batch.hessian_vector(loss, coefs, vs, tag=3)   # the cached-curvature path, as the masked calls below
    active = rng.random(len(dims)) < 0.6
    batch.set_active(active)
    m_values, m_grads = batch.value_and_gradient(loss, coefs)
    m_hvs = batch.hessian_vector(loss, coefs, vs, tag=3)
    elem_active = np.repeat(active, dims)
    assert np.array_equal(m_values, np.where(active, values, 0.0))
    assert np.array_equal(m_grads, np.where(elem_active, grads, 0.0))
    assert np.array_equal(m_hvs, np.where(elem_active, hvs, 0.0))
    # device buffers
    dev = torch.device("cuda", 0)
    with torch.cuda.stream(torch.cuda.ExternalStream(ctx.stream, device=dev)):
        c = torch.from_numpy(coefs).to(dev)
        v = torch.from_numpy(vs).to(dev)
        ov = torch.full((len(dims),), 7.0, dtype=torch.float64, device=dev)
        og = torch.full((batch.d_total,), 7.0, dtype=torch.float64, device=dev)
        oh = torch.full((batch.d_total,), 7.0, dtype=torch.float64, device=dev)
        batch.value_and_gradient_dev(loss, c.data_ptr(), ov.data_ptr(), og.data_ptr())
        batch.hessian_vector_dev(loss, c.data_ptr(), v.data_ptr(), oh.data_ptr(), tag=3)
        ctx.synchronize()
        assert np.array_equal(ov.cpu().numpy(), m_values)
        assert np.array_equal(og.cpu().numpy(), m_grads)
        assert np.array_equal(oh.cpu().numpy(), m_hvs)
    batch.set_active(None)
    again, _ = batch.value_and_gradient(loss, coefs, want_gradient=False)
    assert np.array_equal(again, values)
    with pytest.raises(pml.IllegalArgumentError):
        batch.set_active(np.ones(len(dims) + 1))
    # offsets through the batch handle
    off = rng.normal(scale=0.5, size=n)
    batch.set_offsets(off)
    shifted, _ = batch.value_and_gradient(loss, coefs, want_gradient=False)
    e = int(np.argmax(np.diff(ent_row_ptr)))
    lo, hi = int(ent_row_ptr[e]), int(ent_row_ptr[e + 1])
    sl = slice(int(row_ptr[lo]), int(row_ptr[hi]))
    local = oracle.Data(hi - lo, int(dims[e]), row_ptr[lo:hi + 1] - row_ptr[lo], col[sl], val[sl], label[lo:hi], off[lo:hi])
    a, b = int(batch.coef_ptr[e]), int(batch.coef_ptr[e + 1])
    f_ref, _ = oracle.value_and_gradient(local, loss, coefs[a:b])
    assert abs(shifted[e] - f_ref) <= TOL * abs(f_ref)
    with pytest.raises(pml.IllegalArgumentError):
        batch.set_offsets(off[:-1])
    batch.close()
    ctx.close()


def _entities_with_intercept(rng, n_entities, max_dim, mean_rows):
    """Like _entities, but every entity's last local column is an intercept: 1.0 in every one of its rows."""
    ent_row_ptr, dims, row_ptr, col, val, n = _entities(rng, n_entities, max_dim, mean_rows)
    new_ptr, new_col, new_val = [0], [], []
    for e in range(n_entities):
        for i in range(int(ent_row_ptr[e]), int(ent_row_ptr[e + 1])):
            a, b = int(row_ptr[i]), int(row_ptr[i + 1])
            new_col.extend(col[a:b].tolist() + [int(dims[e])])
            new_val.extend(val[a:b].tolist() + [1.0])
            new_ptr.append(len(new_col))
    return (ent_row_ptr, (dims + 1).astype(np.int32), np.asarray(new_ptr, np.int64), np.asarray(new_col, np.int32),
            np.asarray(new_val, np.float64), n)


@pytest.mark.parametrize("kind", ["SCALE_WITH_STANDARD_DEVIATION", "STANDARDIZATION"])
@pytest.mark.parametrize("loss", [oracle.LOGISTIC, oracle.POISSON])
def test_batch_per_entity_normalization(kind, loss):
    """pml_batch_set_normalization: every entity evaluated in the transformed space of its own context, against the
    oracle run entity by entity with that entity's (factors, shifts, intercept)."""
    import torch
    import photon_ml_b200 as pml
    rng = np.random.default_rng(31 + loss)
    ent_row_ptr, dims, row_ptr, col, val, n = _entities_with_intercept(rng, 250, 10, 14)
    label = util.labels_for(rng, loss, n)
    offset, weight = rng.normal(scale=0.2, size=n), rng.uniform(0.5, 2.0, size=n)
    ctx = pml.Context(0)
    batch = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, col, val, label, offset, weight)
    D = batch.d_total
    factors, shifts, icpt = np.ones(D), np.zeros(D), (dims - 1).astype(np.int32)
    contexts = []
    for e in range(len(dims)):
        lo, hi = int(ent_row_ptr[e]), int(ent_row_ptr[e + 1])
        a = int(batch.coef_ptr[e])
        sl = slice(int(row_ptr[lo]), int(row_ptr[hi]))
        d_e = int(dims[e])
        if hi > lo:
            f, s, _ = util.normalization_context(kind, hi - lo, d_e, row_ptr[lo:hi + 1] - row_ptr[lo], col[sl], val[sl], d_e - 1)
        else:
            f, s = None, None
        f = np.ones(d_e) if f is None else f
        s = np.zeros(d_e) if s is None else s
        factors[a:a + d_e], shifts[a:a + d_e] = f, s
        contexts.append((f, s))
    use_shifts = kind == "STANDARDIZATION"
    batch.set_normalization(factors, shifts if use_shifts else None, icpt)
    coefs = rng.normal(scale=0.2, size=D)
    vs = rng.normal(size=D)
    values, grads = batch.value_and_gradient(loss, coefs)
    hv_cold = batch.hessian_vector(loss, coefs, vs, tag=9)
    hv_cached = batch.hessian_vector(loss, coefs, vs, tag=9)
    dev = torch.device("cuda", 0)
    with torch.cuda.stream(torch.cuda.ExternalStream(ctx.stream, device=dev)):
        c, v = torch.from_numpy(coefs).to(dev), torch.from_numpy(vs).to(dev)
        ov = torch.empty(len(dims), dtype=torch.float64, device=dev)
        og, oh = torch.empty(D, dtype=torch.float64, device=dev), torch.empty(D, dtype=torch.float64, device=dev)
        batch.value_and_gradient_dev(loss, c.data_ptr(), ov.data_ptr(), og.data_ptr())
        batch.hessian_vector_dev(loss, c.data_ptr(), v.data_ptr(), oh.data_ptr(), tag=0)
        ctx.synchronize()
        assert np.array_equal(ov.cpu().numpy(), values) and np.array_equal(og.cpu().numpy(), grads)
        assert np.array_equal(oh.cpu().numpy(), hv_cold)
    for e in range(len(dims)):
        lo, hi = int(ent_row_ptr[e]), int(ent_row_ptr[e + 1])
        a, b = int(batch.coef_ptr[e]), int(batch.coef_ptr[e + 1])
        sl = slice(int(row_ptr[lo]), int(row_ptr[hi]))
        local = oracle.Data(hi - lo, int(dims[e]), row_ptr[lo:hi + 1] - row_ptr[lo], col[sl], val[sl], label[lo:hi],
                            offset[lo:hi], weight[lo:hi])
        f, s = contexts[e]
        s_arg = s if use_shifts else None
        f_ref, g_ref = oracle.value_and_gradient(local, loss, coefs[a:b], f, s_arg, int(dims[e]) - 1)
        hv_ref = oracle.hessian_vector(local, loss, coefs[a:b], vs[a:b], f, s_arg, int(dims[e]) - 1)
        assert abs(values[e] - f_ref) <= 1e-10 * max(abs(f_ref), 1e-300), (e, values[e], f_ref)
        assert util.rel_err(grads[a:b], g_ref) <= 1e-10, e
        assert util.rel_err(hv_cold[a:b], hv_ref) <= 1e-10 and util.rel_err(hv_cached[a:b], hv_ref) <= 1e-10, e
    # the smoothed hinge ignores the context; removing it restores the plain results
    hinge_with, _ = batch.value_and_gradient(oracle.SMOOTHED_HINGE, coefs)
    batch.set_normalization(None, None, None)
    hinge_without, _ = batch.value_and_gradient(oracle.SMOOTHED_HINGE, coefs)
    assert np.array_equal(hinge_with, hinge_without)
    # the requires, entity by entity
    with pytest.raises(pml.IllegalArgumentError):
        batch.set_normalization(None, shifts, None)                       # shift without intercept
    bad = factors.copy()
    bad[int(batch.coef_ptr[3]) + int(dims[3]) - 1] = 2.0
    with pytest.raises(pml.IllegalArgumentError):
        batch.set_normalization(bad, None, icpt)                          # the intercept must not be scaled
    wrong = icpt.copy()
    wrong[:] = 0                                                          # column 0 is not a column of ones
    s2 = shifts.copy()
    s2[batch.coef_ptr[:-1]] = 0.0
    f2 = factors.copy()
    f2[batch.coef_ptr[:-1]] = 1.0
    if use_shifts:
        with pytest.raises(pml.IllegalArgumentError):
            batch.set_normalization(f2, s2, wrong)
    batch.close()
    ctx.close()
