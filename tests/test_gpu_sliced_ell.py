"""The one-lane-per-row margin pass over the sliced-ELL copy of the rows (sell_rows_kernel, pack.cu build_sell) against
the CPU oracle and against the entry-chunked pass it replaces for single-vector passes.  `-m gpu`.

Covered: all three value storages (indices only / fp32 / fp64), coefficients inside and beyond the shared-memory table,
empty rows, a row count that is no multiple of 32, rows of very uneven length (the copy is then not built by default and
built on request), the bank-aware entry order on and off, normalization, and every single-vector pass that runs on it
(value+gradient, cold and cached Hessian-vector, Hessian diagonal, scores).

Reference semantics: the margin x_i . c of ValueAndGradientAggregator.add (lib/function/glm/ValueAndGradientAggregator.scala:
132-153, via LabeledPoint.computeMargin, lib/data/LabeledPoint.scala:62), restated by oracle/glm_oracle.c.
"""
import numpy as np
import pytest

import oracle
from tests import util
from tests.test_gpu_natural_geometry import zipf_rows

pytestmark = pytest.mark.gpu

TOL = 1e-11
KNOBS = ("PML_NO_SELL", "PML_SELL_MAX_OVERHEAD", "PML_SELL_SIGMA_SHIFT", "PML_NO_SELL_BANK_ORDER")


@pytest.fixture(autouse=True)
def no_knobs(monkeypatch):
    for k in KNOBS:
        monkeypatch.delenv(k, raising=False)
    yield


@pytest.fixture(scope="module")
def ctx():
    import photon_ml_b200 as pml
    c = pml.Context(0)
    yield c
    c.close()


def _rows(seed, n, d, draws, empty_every=0):
    row_ptr, col, val = zipf_rows(seed, n, d, draws)
    if empty_every:
        lens = np.diff(row_ptr)
        keep_row = np.ones(n, dtype=bool)
        keep_row[::empty_every] = False
        keep = np.repeat(keep_row, lens)
        lens = np.where(keep_row, lens, 0)
        row_ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        col, val = col[keep], val[keep]
    return row_ptr, col, val


def _check_all(pml, ds, ref, loss, rng, d, norm=None):
    """Every single-vector pass against the oracle, then against the entry-chunked pass on the same data set."""
    w = rng.normal(scale=0.05, size=d)
    v = rng.normal(size=d)
    args = norm or ()
    f_ref, g_ref = oracle.value_and_gradient(ref, loss, w, *args)
    hv_ref = oracle.hessian_vector(ref, loss, w, v, *args)
    diag_ref = oracle.hessian_diagonal(ref, loss, w) if norm is None else None
    out = {}
    for forced in (False, True):
        if forced:
            ds.set_margin_variant(1)        # pins the entry-chunked pass for every row pass
        f, g = ds.value_and_gradient(loss, w)
        assert abs(f - f_ref) <= TOL * abs(f_ref)
        assert util.rel_err(g, g_ref) <= TOL
        # cold: the paired (c, v) pass, also one lane per row; with a tag it fills the curvature cache
        assert util.rel_err(ds.hessian_vector(loss, w, v, tag=9), hv_ref) <= TOL
        hv = ds.hessian_vector(loss, w, v, tag=9)                  # cached: X v is a single-vector pass
        assert util.rel_err(hv, hv_ref) <= TOL
        diag = ds.hessian_diagonal(loss, w)
        if diag_ref is not None:
            assert util.rel_err(diag, diag_ref) <= TOL
        out[forced] = (f, g, hv, diag, ds.scores(w))
    a, b = out[False], out[True]
    assert abs(a[0] - b[0]) <= 1e-13 * abs(b[0])
    for x, y in zip(a[1:], b[1:]):
        assert util.rel_err(x, y) <= 1e-12
    if norm is None:
        # scores are the raw margins x_i . c
        z = np.zeros(ref.n)
        np.add.at(z, np.repeat(np.arange(ref.n), np.diff(ref.row_ptr)), ref.val * w[ref.col])
        assert np.max(np.abs(a[4] - z)) <= 1e-12 * max(1.0, np.max(np.abs(z)))


@pytest.mark.parametrize("storage", ["ones", "f32", "f64"])
@pytest.mark.parametrize("bank_order", [True, False])
def test_sliced_ell_pass_vs_oracle_and_entry_chunked_pass(ctx, monkeypatch, storage, bank_order):
    import photon_ml_b200 as pml
    if not bank_order:
        monkeypatch.setenv("PML_NO_SELL_BANK_ORDER", "1")
    n, d = 70_001, 60_000                      # 70 001 rows: the last slice holds one row; 60 000 columns: beyond the table
    row_ptr, col, val = _rows(31, n, d, 40, empty_every=97)
    rng = np.random.default_rng(32)
    if storage == "ones":
        val = np.ones_like(val)
    elif storage == "f64":
        val = val * rng.uniform(0.9, 1.1, size=len(val))
    label = (rng.random(n) < 0.4).astype(np.float64)
    offset = rng.normal(scale=0.05, size=n)
    weight = rng.uniform(0.5, 1.5, size=n)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label, offset, weight)
    info = ds.info()
    assert info.value_bytes == {"ones": 0, "f32": 4, "f64": 8}[storage]
    assert 0 <= info.sliced_ell_padding_permille <= 150
    ref = oracle.Data(n, d, row_ptr, col, val, label, offset, weight)
    _check_all(pml, ds, ref, oracle.LOGISTIC, rng, d)
    ds.close()


def test_rows_of_uneven_length(ctx, monkeypatch):
    """Power-law row lengths: sorted inside windows the slices are still tight; with tiny windows (sigma 2^5 = one slice,
    i.e. no sorting) padding exceeds the limit and the copy is not built -- unless the limit is lifted."""
    import photon_ml_b200 as pml
    rng = np.random.default_rng(41)
    n, d = 12_011, 5_000
    lens = np.minimum((rng.pareto(1.1, size=n) * 20 + 24).astype(np.int64), 900)
    lens[rng.random(n) < 0.05] = 0
    row_ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    col = np.concatenate([np.sort(rng.choice(d, size=k, replace=False)) for k in lens]).astype(np.int32)
    val = rng.integers(1, 9, size=len(col)) / 4.0
    label = rng.poisson(1.0, size=n).astype(np.float64)
    ref = oracle.Data(n, d, row_ptr, col, val, label)
    seen = []
    for env in ({}, {"PML_SELL_SIGMA_SHIFT": "5"}, {"PML_SELL_SIGMA_SHIFT": "5", "PML_SELL_MAX_OVERHEAD": "100"},
                {"PML_NO_SELL": "1"}):
        for k in KNOBS:
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
        seen.append(ds.info().sliced_ell_padding_permille)
        _check_all(pml, ds, ref, oracle.POISSON, rng, d)
        ds.close()
    assert 0 <= seen[0] <= 150          # sorted by length inside windows of 65 536 rows
    assert seen[1] == -1                # unsorted: too much padding, the entry-chunked stream serves the pass
    assert seen[2] > 150                # built on request all the same
    assert seen[3] == -1


def test_sliced_ell_pass_under_normalization(ctx):
    import photon_ml_b200 as pml
    n, d = 40_000, 30_000
    row_ptr, col, val = _rows(51, n, d, 30)
    icpt = d
    row_ptr, col, val = util.append_intercept(row_ptr, col, val, d)
    rng = np.random.default_rng(52)
    label = (rng.random(n) < 0.5).astype(np.float64)
    ds = pml.Dataset.from_csr(ctx, n, d + 1, row_ptr, col, val, label)
    assert ds.info().sliced_ell_padding_permille >= 0
    factors = rng.uniform(0.5, 2.0, size=d + 1)
    shifts = rng.normal(scale=0.01, size=d + 1)
    factors[icpt], shifts[icpt] = 1.0, 0.0
    ds.set_normalization(factors, shifts, icpt)
    ref = oracle.Data(n, d + 1, row_ptr, col, val, label)
    _check_all(pml, ds, ref, oracle.LOGISTIC, rng, d + 1, norm=(factors, shifts, icpt))
    ds.close()


def test_batches_of_per_entity_problems_keep_the_entry_chunked_pass(ctx):
    """pml_batch data sets mask chunks of the entry stream (pml_batch_set_active): no sliced-ELL copy there."""
    import photon_ml_b200 as pml
    from tests.test_gpu_batch import _entities
    rng = np.random.default_rng(61)
    ent_row_ptr, dims, row_ptr, col, val, n = _entities(rng, 50, 12, 20)
    batch = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, col, val, util.labels_for(rng, oracle.LOGISTIC, n))
    assert batch.dataset.info().sliced_ell_padding_permille == -1
    batch.close()
