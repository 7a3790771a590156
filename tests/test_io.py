"""Packer front-ends (photon_ml_b200.io) against the dense reference implementations in tests/util.py."""
import os

import numpy as np
import pytest

from photon_ml_b200 import io
from tests import util


@pytest.mark.parametrize("name,d", [("heart.txt", 13), ("a9a_1605.libsvm", 123)])
def test_read_libsvm_matches_reference_reader(name, d):
    path = os.path.join(util.GOLDEN, name)
    a = io.read_libsvm(path, feature_dimension=d, add_intercept=True)
    b = util.read_libsvm(path, d=d, add_intercept=True)
    assert a[0] == b[0] and a[1] == b[1]
    for x, y in zip(a[2:], b[2:]):
        assert np.array_equal(x, y)
    c = io.read_libsvm(path, feature_dimension=d, add_intercept=False, max_rows=10)
    assert c[0] == 10 and c[1] == d and c[3].max() < d


@pytest.mark.parametrize("kind", ["NONE", "SCALE_WITH_MAX_MAGNITUDE", "SCALE_WITH_STANDARD_DEVIATION", "STANDARDIZATION"])
def test_summary_and_normalization_factory(kind):
    n, d, row_ptr, col, val, _ = io.read_libsvm(os.path.join(util.GOLDEN, "heart.txt"), feature_dimension=13)
    summary = io.BasicStatisticalSummary(n, d, row_ptr, col, val)
    mean, var, mx, mn = util.column_summary(n, d, row_ptr, col, val)
    assert np.allclose(summary.mean, mean, rtol=1e-13) and np.allclose(summary.variance, var, rtol=1e-11)
    assert np.array_equal(summary.max, mx) and np.array_equal(summary.min, mn)
    ctx = io.normalization_context(kind, summary, d - 1)
    f, s, icpt = util.normalization_context(kind, n, d, row_ptr, col, val, d - 1)
    assert (ctx.factors is None) == (f is None) and (ctx.shifts is None) == (s is None) and ctx.interceptId == icpt
    if f is not None:
        assert np.allclose(ctx.factors, f, rtol=1e-11)
    if s is not None:
        assert np.allclose(ctx.shifts, s, rtol=1e-13)


def test_summary_counts_implicit_zeros():
    # a column stored only on some rows has implicit zeros: min/max/variance must see them
    row_ptr = np.array([0, 1, 1, 2])
    col = np.array([0, 0])
    val = np.array([2.0, 4.0])
    s = io.BasicStatisticalSummary(3, 2, row_ptr, col, val)
    assert s.mean[0] == 2.0 and s.min[0] == 0.0 and s.max[0] == 4.0 and abs(s.variance[0] - 4.0) < 1e-15
    assert s.mean[1] == 0.0 and s.max[1] == 0.0 and s.min[1] == 0.0 and s.variance[1] == 0.0


def test_index_map_projection_per_entity():
    from photon_ml_b200.io import IndexMapProjection
    rng = np.random.default_rng(9)
    n_ent, d = 40, 500
    rows = rng.integers(0, 6, size=n_ent)          # some entities have no rows at all
    ent_row_ptr = np.concatenate([[0], np.cumsum(rows)]).astype(np.int64)
    n = int(ent_row_ptr[-1])
    lens = rng.integers(0, 9, size=n)
    row_ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    col = rng.integers(0, d, size=int(row_ptr[-1])).astype(np.int32)
    proj = IndexMapProjection(ent_row_ptr, row_ptr, col, d)
    assert proj.coef_ptr[-1] == len(proj.original_index) and proj.col_local.shape == col.shape
    for e in range(n_ent):
        a, b = int(row_ptr[ent_row_ptr[e]]), int(row_ptr[ent_row_ptr[e + 1]])
        active = np.unique(col[a:b])
        lo, hi = int(proj.coef_ptr[e]), int(proj.coef_ptr[e + 1])
        assert proj.ent_dim[e] == len(active) == hi - lo
        assert np.array_equal(proj.original_index[lo:hi], active)                 # one-to-one, ascending
        assert np.array_equal(proj.original_index[lo + proj.col_local[a:b]], col[a:b])   # every entry maps back
        assert (proj.col_local[a:b] >= 0).all() and (proj.col_local[a:b] < max(1, hi - lo)).all()
    w = rng.normal(size=d)
    e = int(np.argmax(proj.ent_dim))
    local = proj.project_coefficients(w, e)
    concat = np.zeros(int(proj.coef_ptr[-1]))
    concat[int(proj.coef_ptr[e]):int(proj.coef_ptr[e + 1])] = local
    back = proj.to_original_space(concat, e)
    active = proj.original_index[int(proj.coef_ptr[e]):int(proj.coef_ptr[e + 1])]
    assert np.array_equal(back[active], w[active]) and np.count_nonzero(back) <= len(active)
    with pytest.raises(ValueError):
        IndexMapProjection(ent_row_ptr, row_ptr, col, int(col.max()))
