"""The CUDA path against the CPU oracle in the geometry the packer picks BY ITSELF.  `-m gpu`.

tests/test_gpu_parity.py forces 37-row tiles and 101-row super blocks onto small inputs; that exercises every
code path of the tiled layout except the ones that only large inputs reach: (tile, column) runs longer than a
256-entry chunk (the no-scan path of seg_spmv_tiled_kernel and the bank-aware order of pack.cu), the 96 KB
coefficient table of the CSR pass, tiles of thousands of rows, and the default super blocks (above 6 M rows).  Here the inputs are
bench-shaped (Zipf(1.1) columns, ~50 entries per row, values in {1,2,3,4}/4) and no packing knob is set.

Reference semantics: ValueAndGradientAggregator.scala:132-153,199-221; HessianVectorAggregator.scala:96-121;
HessianDiagonalAggregator.scala:51-61 (restated by oracle/glm_oracle.c).
"""
import os

import numpy as np
import pytest

import oracle
from tests import util

pytestmark = pytest.mark.gpu

TOL = 1e-11
KNOBS = ("PML_TILE_MIN_ROWS", "PML_TILE_ROWS", "PML_NO_TILING", "PML_TILE_MAX_ROWS", "PML_TILE_PARTIAL_MB",
         "PML_TILE_MIN_RUN", "PML_SUPER_ROWS", "PML_SUPER_MIN_ROWS", "PML_NO_BANK_ORDER", "PML_NO_UNIFORM_FAST")


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


def zipf_rows(seed, n, d, draws, a=1.1, sort_rows=True):
    """Bench-shaped rows: `draws` Zipf(a) columns per row, duplicates merged, values in {1,2,3,4}/4."""
    rng = np.random.default_rng(seed)
    p = np.arange(1, d + 1, dtype=np.float64) ** -a
    cdf = np.cumsum(p / p.sum())
    cols = np.minimum(np.searchsorted(cdf, rng.random((n, draws))), d - 1).astype(np.int32)
    cols.sort(axis=1)
    keep = np.ones_like(cols, dtype=bool)
    keep[:, 1:] = cols[:, 1:] != cols[:, :-1]
    vals = rng.integers(1, 5, size=(n, draws)).astype(np.float64) / 4.0
    row_ptr = np.concatenate([[0], np.cumsum(keep.sum(axis=1))]).astype(np.int64)
    # a random permutation of the column ids, so that "hot" is something the packer has to find
    perm = rng.permutation(d).astype(np.int32)
    col = perm[cols[keep]]
    val = vals[keep]
    if not sort_rows:   # Breeze SparseVectors need not be sorted (lib/model/Coefficients.scala:26-27)
        return row_ptr, col, val
    # per-row ascending column order (what VectorUtils.toSparseVector produces)
    order = np.lexsort((col, np.repeat(np.arange(n), np.diff(row_ptr))))
    return row_ptr, col[order], val[order]


@pytest.fixture(scope="module")
def natural(ctx):
    n, d = 320_000, 200_000
    row_ptr, col, val = zipf_rows(11, n, d, 66)
    rng = np.random.default_rng(12)
    w_true = rng.normal(scale=0.1, size=d) * (rng.random(d) < 0.02)
    z = np.zeros(n)
    np.add.at(z, np.repeat(np.arange(n), np.diff(row_ptr)), val * w_true[col])
    return n, d, row_ptr, col, val, np.clip(z, -4, 4), rng


def _labels(loss, z, rng):
    if loss == oracle.POISSON:
        return rng.poisson(np.exp(z)).astype(np.float64)
    if loss == oracle.SQUARED:
        return z + rng.normal(scale=0.1, size=len(z))
    return (rng.random(len(z)) < 1.0 / (1.0 + np.exp(-z))).astype(np.float64)


@pytest.mark.parametrize("loss", [oracle.LOGISTIC, oracle.POISSON, oracle.SQUARED, oracle.SMOOTHED_HINGE])
def test_bench_shaped_rows_default_tiling_vs_oracle(ctx, natural, loss):
    import photon_ml_b200 as pml
    n, d, row_ptr, col, val, z, rng = natural
    label = _labels(loss, z, rng)
    offset = rng.normal(scale=0.05, size=n)
    weight = rng.uniform(0.5, 1.5, size=n)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label, offset, weight)
    info = ds.info()
    # the geometry under test is the packer's own: tiles of thousands of rows, thousands of hot columns
    assert info.value_bytes == 4
    assert info.tiled_tiles >= 148 and info.tiled_tile_rows > 1000 and info.tiled_hot_cols > 1000
    ref = oracle.Data(n, d, row_ptr, col, val, label, offset, weight)
    cores = oracle.max_threads()
    for scale in (0.02, 0.2):
        w = rng.normal(scale=scale, size=d)
        v = rng.normal(size=d)
        f_ref, g_ref = oracle.value_and_gradient(ref, loss, w, parts=cores)
        f, g = ds.value_and_gradient(loss, w)
        assert abs(f - f_ref) <= TOL * abs(f_ref)
        assert util.rel_err(g, g_ref) <= TOL
        assert np.max(np.abs(g - g_ref)) <= TOL * np.max(np.abs(g_ref))
        if loss == oracle.SMOOTHED_HINGE:
            continue
        hv_ref = oracle.hessian_vector(ref, loss, w, v)
        assert util.rel_err(ds.hessian_vector(loss, w, v), hv_ref) <= TOL
        assert util.rel_err(ds.hessian_vector(loss, w, v, tag=5), hv_ref) <= TOL
        assert util.rel_err(ds.hessian_vector(loss, w, v, tag=5), hv_ref) <= TOL      # cached curvature
        assert util.rel_err(ds.hessian_diagonal(loss, w), oracle.hessian_diagonal(ref, loss, w)) <= TOL
    ds.close()


def test_bench_shaped_rows_with_normalization_and_indicator_values(ctx, natural):
    """Same rows with every value 1.0 (indices-only storage) under a STANDARDIZATION-style context."""
    import photon_ml_b200 as pml
    n, d, row_ptr, col, val, z, rng = natural
    ones = np.ones_like(val)
    icpt = d
    row_ptr2, col2, val2 = util.append_intercept(row_ptr, col, ones, d)
    label = _labels(oracle.LOGISTIC, z, rng)
    ds = pml.Dataset.from_csr(ctx, n, d + 1, row_ptr2, col2, val2, label)
    assert ds.info().value_bytes == 0 and ds.info().tiled_hot_cols > 1000
    factors = rng.uniform(0.5, 2.0, size=d + 1)
    shifts = rng.normal(scale=0.01, size=d + 1)
    factors[icpt], shifts[icpt] = 1.0, 0.0
    ds.set_normalization(factors, shifts, icpt)
    ref = oracle.Data(n, d + 1, row_ptr2, col2, val2, label)
    w = rng.normal(scale=0.05, size=d + 1)
    v = rng.normal(size=d + 1)
    f_ref, g_ref = oracle.value_and_gradient(ref, oracle.LOGISTIC, w, factors, shifts, icpt)
    f, g = ds.value_and_gradient(pml.LOGISTIC, w)
    assert abs(f - f_ref) <= TOL * abs(f_ref)
    assert util.rel_err(g, g_ref) <= TOL
    hv_ref = oracle.hessian_vector(ref, oracle.LOGISTIC, w, v, factors, shifts, icpt)
    assert util.rel_err(ds.hessian_vector(pml.LOGISTIC, w, v), hv_ref) <= TOL
    ds.close()


def test_more_than_twelve_million_rows_super_blocks_vs_oracle(ctx):
    """Above 6 M rows the cold columns are cut into super blocks of <= 6 M rows by default (pack.cu): three of them
    here; the oracle streams the same 12.7 M rows (partitioned over the host cores)."""
    import photon_ml_b200 as pml
    n, d = 12_700_000, 300_000
    row_ptr, col, val = zipf_rows(21, n, d, 6, sort_rows=False)
    rng = np.random.default_rng(22)
    label = (rng.random(n) < 0.3).astype(np.float64)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
    info = ds.info()
    assert info.tiled_tiles >= 148 and info.tiled_hot_cols >= 16
    ref = oracle.Data(n, d, row_ptr, col, val, label)
    w = rng.normal(scale=0.1, size=d)
    f_ref, g_ref = oracle.value_and_gradient(ref, oracle.LOGISTIC, w, parts=oracle.max_threads())
    f, g = ds.value_and_gradient(pml.LOGISTIC, w)
    assert abs(f - f_ref) <= TOL * abs(f_ref)
    assert util.rel_err(g, g_ref) <= TOL
    # bit-reproducible, and the packing switches change the order of the sums only
    f2, g2 = ds.value_and_gradient(pml.LOGISTIC, w)
    assert f2 == f and np.array_equal(g, g2)
    ds.close()


@pytest.mark.parametrize("knob", ["PML_NO_BANK_ORDER", "PML_NO_UNIFORM_FAST"])
def test_packing_switches_do_not_change_results(ctx, natural, knob, monkeypatch):
    """The bank-aware order and the no-scan path only reorder terms inside sums: results agree to rounding."""
    import photon_ml_b200 as pml
    n, d, row_ptr, col, val, z, rng = natural
    label = _labels(oracle.POISSON, z, rng)
    w = rng.normal(scale=0.05, size=d)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
    f0, g0 = ds.value_and_gradient(pml.POISSON, w)
    ds.close()
    monkeypatch.setenv(knob, "1")
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
    f1, g1 = ds.value_and_gradient(pml.POISSON, w)
    ds.close()
    assert abs(f0 - f1) <= 1e-13 * abs(f0)
    assert util.rel_err(g1, g0) <= 1e-13
