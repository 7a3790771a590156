"""The BASELINE.json configurations as parity cases: configs[2..4] at reduced size against the
oracle, and configs[1] at FULL size (10 M rows x 1 M features, 5e8 entries) through size-independent
properties (linearity of the Hessian-vector product, gradient additivity over row shards, the
directional derivative of the value, a checksum of per-row scores).  `-m gpu`."""
import numpy as np
import pytest

import oracle
from tests import util

pytestmark = pytest.mark.gpu
TOL = 1e-11


@pytest.fixture(scope="module")
def ctx():
    import photon_ml_b200 as pml
    c = pml.Context(0)
    yield c
    c.close()


def test_config2_criteo_shaped_logistic_reduced(ctx):
    # configs[2]: 39 entries/row = 13 "numeric" columns from a small dense range + 26 hashed Zipf(1.05)
    # categorical columns, values 1.0, CTR ~3 %; reduced to 40 k rows x 200 k hashed features
    import photon_ml_b200 as pml
    rng = np.random.default_rng(3)
    n, d = 40_000, 200_000
    dense = rng.integers(0, 64, size=(n, 13))
    p = 1.0 / np.arange(1, d - 64 + 1) ** 1.05
    cdf = np.cumsum(p / p.sum())
    hashed = 64 + np.searchsorted(cdf, rng.random((n, 26))).clip(0, d - 65)
    col = np.sort(np.concatenate([dense, hashed], axis=1), axis=1).astype(np.int32).ravel()   # duplicates kept
    row_ptr = np.arange(n + 1, dtype=np.int64) * 39
    val = np.ones(n * 39)
    label = (rng.random(n) < 0.03).astype(np.float64)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
    ref = oracle.Data(n, d, row_ptr, col, val, label)
    assert ds.info().value_bytes == 0     # every value is 1.0: only the indices are stored
    w = rng.normal(scale=0.05, size=d)
    f_ref, g_ref = oracle.value_and_gradient(ref, oracle.LOGISTIC, w, parts=4)
    f, g = ds.value_and_gradient(oracle.LOGISTIC, w)
    assert abs(f - f_ref) <= TOL * abs(f_ref) and util.rel_err(g, g_ref) <= TOL
    ds.close()


def test_config3_glmix_shaped_fixed_plus_random_effects(ctx):
    # configs[3] in miniature: fixed effect + per-user random effect, one coordinate-descent pass:
    # FE scores become the RE offsets (Coordinate.scala:59-62), RE solved as a batch
    import photon_ml_b200 as pml
    rng = np.random.default_rng(4)
    n_users, d_fe, d_re = 300, 500, 21
    rows_per_user = np.maximum(rng.lognormal(3.0, 0.8, size=n_users).astype(np.int64), 5)
    n = int(rows_per_user.sum())
    fe_ptr, fe_col, fe_val = util.random_csr(rng, n, d_fe, 30)
    re_ptr, re_col, re_val = util.random_csr(rng, n, d_re, 8)
    label = (rng.random(n) < 0.4).astype(np.float64)
    fe = pml.Dataset.from_csr(ctx, n, d_fe, fe_ptr, fe_col, fe_val, label)
    w_fe = rng.normal(scale=0.1, size=d_fe)
    scores = fe.scores(w_fe)                                  # FixedEffectCoordinate.score
    ent_row_ptr = np.concatenate([[0], np.cumsum(rows_per_user)]).astype(np.int64)
    batch = pml.Batch(ctx, ent_row_ptr, np.full(n_users, d_re, np.int32), re_ptr, re_col, re_val, label)
    batch.dataset.set_offsets(scores)                          # residual from the other coordinate
    coefs = rng.normal(scale=0.2, size=batch.d_total)
    values, grads = batch.value_and_gradient(oracle.LOGISTIC, coefs)
    X_fe = util.csr_to_dense(n, d_fe, fe_ptr, fe_col, fe_val)
    assert util.rel_err(scores, X_fe @ w_fe) <= TOL
    for e in (0, int(np.argmax(rows_per_user)), n_users - 1):
        lo, hi = int(ent_row_ptr[e]), int(ent_row_ptr[e + 1])
        sl = slice(int(re_ptr[lo]), int(re_ptr[hi]))
        local = oracle.Data(hi - lo, d_re, re_ptr[lo:hi + 1] - re_ptr[lo], re_col[sl], re_val[sl], label[lo:hi], scores[lo:hi])
        f_ref, g_ref = oracle.value_and_gradient(local, oracle.LOGISTIC, coefs[e * d_re:(e + 1) * d_re])
        assert abs(values[e] - f_ref) <= TOL * abs(f_ref)
        assert util.rel_err(grads[e * d_re:(e + 1) * d_re], g_ref) <= TOL
    batch.close()
    fe.close()


def test_config4_game_ads_ctr_shaped_many_tiny_entities(ctx):
    # configs[4] in miniature: power-law rows per entity (mean ~4), ~10 entries/row, d_e <= 32
    import photon_ml_b200 as pml
    rng = np.random.default_rng(5)
    n_ent = 20_000
    rows = np.minimum(rng.zipf(1.8, size=n_ent), 400).astype(np.int64)
    dims = rng.integers(4, 33, size=n_ent).astype(np.int32)
    ent_row_ptr = np.concatenate([[0], np.cumsum(rows)]).astype(np.int64)
    n = int(ent_row_ptr[-1])
    ent_of_row = np.repeat(np.arange(n_ent), rows)
    k = 10
    col = (rng.random((n, k)) * dims[ent_of_row][:, None]).astype(np.int32)
    col = np.sort(col, axis=1).ravel()
    row_ptr = np.arange(n + 1, dtype=np.int64) * k
    val = np.ones(n * k)
    label = (rng.random(n) < 0.1).astype(np.float64)
    batch = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, col, val, label)
    coefs = rng.normal(scale=0.3, size=batch.d_total)
    values, grads = batch.value_and_gradient(oracle.LOGISTIC, coefs)
    # the whole batch equals one block-diagonal problem for the oracle
    gcol = (col.reshape(n, k) + batch.coef_ptr[ent_of_row][:, None]).astype(np.int32).ravel()
    ref = oracle.Data(n, batch.d_total, row_ptr, gcol, val, label)
    f_ref, g_ref = oracle.value_and_gradient(ref, oracle.LOGISTIC, coefs)
    assert abs(values.sum() - f_ref) <= 1e-10 * abs(f_ref)
    assert util.rel_err(grads, g_ref) <= TOL
    batch.close()


def test_config1_full_size_properties(ctx):
    """configs[1] at full size: 10 M rows x 1 M features, ~5e8 entries (the bench workload)."""
    import torch
    import bench
    import photon_ml_b200 as pml
    dev = torch.device("cuda", 0)
    chunks = list(range(bench.N_ROWS // bench.GEN_CHUNK))
    row_ptr, cols, vals, labels = bench.generate_rows(torch, dev, chunks)
    n = labels.numel()
    d = bench.N_FEATS
    full = pml.Dataset.from_device_pointers(ctx, n, d, row_ptr.data_ptr(), cols.data_ptr(), vals.data_ptr(), pml.abi.F32,
                                            labels.data_ptr(), storage=pml.STORE_F32)
    assert full.info().nnz > 4.9e8
    # two row shards built from the same arrays: raw sums must add up to the full problem's
    cut = n // 3
    nnz_cut = int(row_ptr[cut].item())
    rp_b = (row_ptr[cut:] - nnz_cut).contiguous()
    a = pml.Dataset.from_device_pointers(ctx, cut, d, row_ptr.data_ptr(), cols.data_ptr(), vals.data_ptr(), pml.abi.F32,
                                         labels.data_ptr(), storage=pml.STORE_F32)
    b = pml.Dataset.from_device_pointers(ctx, n - cut, d, rp_b.data_ptr(), cols[nnz_cut:].data_ptr(),
                                         vals[nnz_cut:].data_ptr(), pml.abi.F32, labels[cut:].data_ptr(),
                                         storage=pml.STORE_F32)
    del row_ptr, cols, vals
    rng = np.random.default_rng(9)
    w = rng.normal(scale=0.02, size=d)
    v1, v2 = rng.normal(size=d), rng.normal(size=d)
    f, g = full.value_and_gradient(pml.POISSON, w)
    fa, ga = a.value_and_gradient(pml.POISSON, w)
    fb, gb = b.value_and_gradient(pml.POISSON, w)
    assert abs(f - (fa + fb)) <= 1e-12 * abs(f)                       # additivity over row shards (merge)
    assert util.rel_err(g, ga + gb) <= 1e-12
    h1 = full.hessian_vector(pml.POISSON, w, v1, tag=1)
    h2 = full.hessian_vector(pml.POISSON, w, v2, tag=1)
    h12 = full.hessian_vector(pml.POISSON, w, 0.5 * v1 - 2.0 * v2, tag=1)
    assert util.rel_err(h12, 0.5 * h1 - 2.0 * h2) <= 1e-11                    # linearity in v
    assert util.rel_err(full.hessian_vector(pml.POISSON, w, v1), h1) <= 1e-13   # cached == cold
    assert abs(v1.dot(h2) - v2.dot(h1)) <= 1e-10 * abs(v1.dot(h2))            # symmetry of the Hessian
    eps = 1e-6
    fp, _ = full.value_and_gradient(pml.POISSON, w + eps * v1, want_gradient=False)
    fm, _ = full.value_and_gradient(pml.POISSON, w - eps * v1, want_gradient=False)
    assert abs((fp - fm) / (2 * eps) - g.dot(v1)) <= 1e-6 * abs(g.dot(v1))    # directional derivative
    # scores: checksum against the gradient identity  sum_i r_i * score_i(v) == g . v  with r = dl/dz
    sq = full.value_and_gradient(pml.SQUARED, w)                               # squared loss: r_i = z_i - y_i
    scores = full.scores(w)
    y = labels.cpu().numpy()
    assert abs(0.5 * np.sum((scores - y) ** 2) - sq[0]) <= 1e-11 * sq[0]
    for ds in (full, a, b):
        ds.close()


def test_coordinate_descent_glue_stays_on_the_device(ctx):
    # One GAME pass in miniature without host round trips of per-row data: fixed-effect scores
    # (FixedEffectCoordinate.score) become the random-effect coordinate's offsets
    # (DataSet.addScoresToOffsets) through a row map, because the random-effect rows are grouped by entity.
    import torch
    import photon_ml_b200 as pml
    rng = np.random.default_rng(6)
    n, d_fe, d_re, n_ent = 5000, 300, 12, 200
    fe_ptr, fe_col, fe_val = util.random_csr(rng, n, d_fe, 15)
    label = (rng.random(n) < 0.3).astype(np.float64)
    base_offset = rng.normal(scale=0.1, size=n)
    entity_of_row = rng.integers(0, n_ent, size=n)
    row_map = np.argsort(entity_of_row, kind="stable").astype(np.int64)     # RE row i is FE row row_map[i]
    rows = np.bincount(entity_of_row, minlength=n_ent)
    ent_row_ptr = np.concatenate([[0], np.cumsum(rows)]).astype(np.int64)
    re_ptr, re_col, re_val = util.random_csr(rng, n, d_re, 5)                # already in RE (entity-grouped) order
    fe = pml.Dataset.from_csr(ctx, n, d_fe, fe_ptr, fe_col, fe_val, label, base_offset)
    batch = pml.Batch(ctx, ent_row_ptr, np.full(n_ent, d_re, np.int32), re_ptr, re_col, re_val, label[row_map])
    dev = torch.device("cuda", 0)
    w_fe = torch.tensor(rng.normal(scale=0.2, size=d_fe), device=dev)
    t_scores = torch.empty(n, dtype=torch.float64, device=dev)
    t_base = torch.tensor(base_offset[row_map], device=dev)
    t_map = torch.tensor(row_map, device=dev)
    t_off = torch.empty(n, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    fe.scores_dev(w_fe.data_ptr(), t_scores.data_ptr())
    pml.abi.offsets_from_scores_dev(ctx, n, t_base.data_ptr(), t_scores.data_ptr(), t_map.data_ptr(), t_off.data_ptr())
    batch.dataset.set_offsets_dev(t_off.data_ptr())
    coefs = rng.normal(scale=0.3, size=batch.d_total)
    values, grads = batch.value_and_gradient(pml.LOGISTIC, coefs)
    X_fe = util.csr_to_dense(n, d_fe, fe_ptr, fe_col, fe_val)
    off_ref = base_offset[row_map] + (X_fe @ w_fe.cpu().numpy())[row_map]
    assert util.rel_err(t_off.cpu().numpy(), off_ref) <= TOL
    gcol = (re_col + np.repeat(batch.coef_ptr[:-1], rows)[np.repeat(np.arange(n), np.diff(re_ptr))]).astype(np.int32)
    ref = oracle.Data(n, batch.d_total, re_ptr, gcol, re_val, label[row_map], off_ref)
    f_ref, g_ref = oracle.value_and_gradient(ref, oracle.LOGISTIC, coefs)
    assert abs(values.sum() - f_ref) <= 1e-10 * abs(f_ref) and util.rel_err(grads, g_ref) <= TOL
    # training-loss evaluator = value only: same number, no column pass
    launches0 = ctx.launch_count
    f_only, _ = batch.dataset.value_and_gradient(pml.LOGISTIC, coefs, want_gradient=False)
    value_only_launches = ctx.launch_count - launches0
    launches0 = ctx.launch_count
    f_full, _ = batch.dataset.value_and_gradient(pml.LOGISTIC, coefs)
    assert f_only == f_full and value_only_launches < ctx.launch_count - launches0
    batch.dataset.set_offsets_dev(None)
    batch.close()
    fe.close()
