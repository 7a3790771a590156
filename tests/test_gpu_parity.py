"""Parity of the CUDA path (through the C ABI) against the CPU oracle.  `-m gpu`.

Bar (BASELINE.json north star): relative error <= 1e-9 on objective value, gradient and
Hessian-vector product per evaluation, norm-wise.  The kernels accumulate in fp64 in
a different (fixed) order than the reference's sequential loop, so differences are ~1e-15; the
tests assert TOL = 1e-11 to leave the north-star bar two orders of head-room.
"""
import os

import numpy as np
import pytest

import oracle
from tests import util

pytestmark = pytest.mark.gpu

TOL = 1e-11
LOSSES = [oracle.LOGISTIC, oracle.POISSON, oracle.SQUARED]
VARIANTS = [0, 1, 2, 4, 8, 16, 32]


@pytest.fixture(autouse=True, params=["plain-csc", "tiled-csc", "tiled-csc-super"])
def csc_layout(request, monkeypatch):
    """Every test runs with the plain CSC layout, with the row-tiled one (hot columns gathered from
    shared-memory slices of r), forced onto these small inputs with tiny tiles, and with the cold columns
    additionally cut into super blocks of rows (per-(super block, column) sums reduced afterwards)."""
    if request.param.startswith("tiled-csc"):
        monkeypatch.setenv("PML_TILE_MIN_ROWS", "0")
        monkeypatch.setenv("PML_TILE_ROWS", "37")
        monkeypatch.setenv("PML_SUPER_ROWS", "101" if request.param.endswith("super") else "1000000000")
    else:
        monkeypatch.setenv("PML_NO_TILING", "1")
    yield


@pytest.fixture(scope="module")
def ctx():
    import photon_ml_b200 as pml
    c = pml.Context(0)
    yield c
    c.close()


def _pair(ctx, n, d, row_ptr, col, val, label, offset=None, weight=None, storage=None):
    import photon_ml_b200 as pml
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label, offset, weight,
                              storage=pml.STORE_AUTO if storage is None else storage)
    ref = oracle.Data(n, d, row_ptr, col, val, label, offset, weight)
    return ds, ref


def _check_all(ds, ref, loss, rng, factors=None, shifts=None, intercept=None, scale=0.3, draws=3):
    for _ in range(draws):
        w = rng.normal(scale=scale, size=ref.d)
        v = rng.normal(size=ref.d)
        f_ref, g_ref = oracle.value_and_gradient(ref, loss, w, factors, shifts, intercept)
        f, g = ds.value_and_gradient(loss, w)
        assert abs(f - f_ref) <= TOL * max(abs(f_ref), 1e-300), (f, f_ref)
        assert util.rel_err(g, g_ref) <= TOL
        f_only, none = ds.value_and_gradient(loss, w, want_gradient=False)
        assert f_only == f and none is None
        if loss == oracle.SMOOTHED_HINGE:
            continue
        hv_ref = oracle.hessian_vector(ref, loss, w, v, factors, shifts, intercept)
        assert util.rel_err(ds.hessian_vector(loss, w, v), hv_ref) <= TOL
        # tagged: first call fills the cache, the second reuses it with another direction
        assert util.rel_err(ds.hessian_vector(loss, w, v, tag=77), hv_ref) <= TOL
        v2 = rng.normal(size=ref.d)
        hv2_ref = oracle.hessian_vector(ref, loss, w, v2, factors, shifts, intercept)
        assert util.rel_err(ds.hessian_vector(loss, None if False else w, v2, tag=77), hv2_ref) <= TOL
        # a new coefficient vector under a new tag must not see the old cache
        w3 = w + 0.1
        hv3_ref = oracle.hessian_vector(ref, loss, w3, v2, factors, shifts, intercept)
        assert util.rel_err(ds.hessian_vector(loss, w3, v2, tag=78), hv3_ref) <= TOL
        diag_ref = oracle.hessian_diagonal(ref, loss, w)
        assert util.rel_err(ds.hessian_diagonal(loss, w), diag_ref) <= TOL


def _a9a():
    return util.read_libsvm(os.path.join(util.GOLDEN, "a9a_1605.libsvm"), d=123, add_intercept=True)


@pytest.mark.parametrize("variant", VARIANTS)
def test_a1a_shaped_logistic_all_row_variants(ctx, variant):
    # config 1 of BASELINE.json: the first 1605 rows of the reference's a9a fixture + intercept
    n, d, row_ptr, col, val, label = _a9a()
    ds, ref = _pair(ctx, n, d, row_ptr, col, val, label)
    info = ds.info()
    assert info.value_bytes == 0  # all values are exactly 1.0: AUTO stores the indices only
    if os.environ.get("PML_TILE_ROWS"):   # the tiled layout really is in use in the tiled-csc runs
        assert info.tiled_hot_cols >= 16 and info.tiled_tiles == (n + 36) // 37 and info.tiled_tile_rows == 37
    else:
        assert info.tiled_hot_cols == 0
    ds.set_margin_variant(variant)
    _check_all(ds, ref, oracle.LOGISTIC, np.random.default_rng(1))
    ds.close()


@pytest.mark.parametrize("loss", LOSSES + [oracle.SMOOTHED_HINGE])
def test_a1a_shaped_every_loss(ctx, loss):
    n, d, row_ptr, col, val, label = _a9a()
    lab = label if loss != oracle.POISSON else label * 2.0
    ds, ref = _pair(ctx, n, d, row_ptr, col, val, lab)
    _check_all(ds, ref, loss, np.random.default_rng(2), scale=0.2)
    ds.close()


@pytest.mark.parametrize("loss", LOSSES)
@pytest.mark.parametrize("kind", ["SCALE_WITH_MAX_MAGNITUDE", "SCALE_WITH_STANDARD_DEVIATION", "STANDARDIZATION"])
def test_heart_with_normalization_fp64_values(ctx, loss, kind):
    # heart.txt holds 2.4, 0.3, ... which fp32 cannot represent: AUTO must pick 8-byte storage
    n, d, row_ptr, col, val, label = util.read_libsvm(os.path.join(util.GOLDEN, "heart.txt"), d=13, add_intercept=True)
    lab = label if loss != oracle.POISSON else label * 3.0
    ds, ref = _pair(ctx, n, d, row_ptr, col, val, lab)
    assert ds.info().value_bytes == 8
    f, s, icpt = util.normalization_context(kind, n, d, row_ptr, col, val, d - 1)
    ds.set_normalization(f, s, icpt)
    _check_all(ds, ref, loss, np.random.default_rng(3), f, s, icpt, scale=0.05 if loss == oracle.POISSON else 0.3)
    # the smoothed hinge ignores the context (SmoothedHingeLossFunction.scala:73-84)
    w = np.random.default_rng(4).normal(scale=0.01, size=d)
    fh_ref, gh_ref = oracle.value_and_gradient(ref, oracle.SMOOTHED_HINGE, w)
    fh, gh = ds.value_and_gradient(oracle.SMOOTHED_HINGE, w)
    assert abs(fh - fh_ref) <= TOL * abs(fh_ref) and util.rel_err(gh, gh_ref) <= TOL
    ds.close()


@pytest.mark.parametrize("variant", [1, 8])
def test_indicator_values_stored_or_implicit_give_the_same_bits(ctx, variant):
    """PML_STORE_ONES (what AUTO picks for a9a) against the same data with its 1.0s stored as fp32: multiplying by
    1.0 is exact, so value, gradient, Hessian-vector product and diagonal must agree to the last bit."""
    import photon_ml_b200 as pml
    n, d, row_ptr, col, val, label = _a9a()
    rng = np.random.default_rng(8)
    offset, weight = rng.normal(scale=0.2, size=n), rng.uniform(0.5, 2.0, size=n)
    f, s, icpt = util.normalization_context("STANDARDIZATION", n, d, row_ptr, col, val, d - 1)
    w, v = rng.normal(scale=0.2, size=d), rng.normal(size=d)
    got = []
    for storage in (pml.STORE_ONES, pml.STORE_F32, pml.STORE_F64):
        ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label, offset, weight, storage=storage)
        assert ds.info().value_bytes == {pml.STORE_ONES: 0, pml.STORE_F32: 4, pml.STORE_F64: 8}[storage]
        ds.set_margin_variant(variant)
        ds.set_normalization(f, s, icpt)
        fv, g = ds.value_and_gradient(oracle.LOGISTIC, w)
        got.append((fv, g, ds.hessian_vector(oracle.LOGISTIC, w, v), ds.hessian_vector(oracle.LOGISTIC, w, v, tag=5),
                    ds.hessian_vector(oracle.LOGISTIC, w, v, tag=5), ds.hessian_diagonal(oracle.LOGISTIC, w), ds.scores(w)))
        ds.close()
    for other in got[1:]:
        assert got[0][0] == other[0]
        for a, b in zip(got[0][1:], other[1:]):
            assert np.array_equal(a, b)
    ref = oracle.Data(n, d, row_ptr, col, val, label, offset, weight)
    f_ref, g_ref = oracle.value_and_gradient(ref, oracle.LOGISTIC, w, f, s, icpt)
    assert abs(got[0][0] - f_ref) <= TOL * abs(f_ref) and util.rel_err(got[0][1], g_ref) <= TOL
    # values other than 1.0 are refused in this mode
    val2 = val.copy()
    val2[7] = 0.5
    with pytest.raises(pml.IllegalArgumentError):
        pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val2, label, storage=pml.STORE_ONES)


def test_heart_fp32_storage_is_refused_but_f32_input_is_fine(ctx):
    import photon_ml_b200 as pml
    n, d, row_ptr, col, val, label = util.read_libsvm(os.path.join(util.GOLDEN, "heart.txt"), d=13, add_intercept=True)
    with pytest.raises(pml.IllegalArgumentError):
        pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label, storage=pml.STORE_F32)
    # parity is defined on the values as packed: feed fp32 values, compare with the oracle on the widened ones
    v32 = val.astype(np.float32)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, v32, label)
    assert ds.info().value_bytes == 4
    ref = oracle.Data(n, d, row_ptr, col, v32.astype(np.float64), label)
    _check_all(ds, ref, oracle.LOGISTIC, np.random.default_rng(5), scale=0.01)
    ds.close()


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("loss", [oracle.LOGISTIC, oracle.POISSON])
def test_ragged_skewed_with_weights_offsets(ctx, variant, loss):
    # hot columns span several 4096-entry chunks (fix-up path), 5 % empty rows, many empty columns
    rng = np.random.default_rng(10 + variant)
    n, d = 6000, 4000
    row_ptr, col, val = util.random_csr(rng, n, d, 9, skew=1.3, empty_rows=0.05)
    off = rng.normal(scale=0.2, size=n)
    wt = rng.uniform(0.2, 3.0, size=n)
    label = util.labels_for(rng, loss, n)
    ds, ref = _pair(ctx, n, d, row_ptr, col, val, label, off, wt)
    info = ds.info()
    assert info.n_chunks > 5 and info.nonempty_cols < d
    ds.set_margin_variant(variant)
    _check_all(ds, ref, loss, rng, scale=0.1, draws=2)
    ds.close()


def test_long_rows_and_duplicate_columns(ctx):
    rng = np.random.default_rng(20)
    n, d = 300, 5000
    row_ptr, col, val = util.random_csr(rng, n, d, 600, distinct=False)  # rows longer than the staging buffer
    label = util.labels_for(rng, 0, n)
    ds, ref = _pair(ctx, n, d, row_ptr, col, val, label)
    for variant in VARIANTS:
        ds.set_margin_variant(variant)
        _check_all(ds, ref, oracle.LOGISTIC, rng, scale=0.02, draws=1)
    ds.close()


@pytest.mark.parametrize("shape", [(1, 1, 1), (1, 7, 3), (40, 1, 1), (33, 5, 0), (20000, 1, 1), (20000, 3, 2)])
def test_degenerate_shapes(ctx, shape):
    n, d, k = shape
    rng = np.random.default_rng(30)
    row_ptr, col, val = util.random_csr(rng, n, d, k) if k > 0 else (np.zeros(n + 1, np.int64), np.zeros(0, np.int32), np.zeros(0))
    label = util.labels_for(rng, 0, n)
    ds, ref = _pair(ctx, n, d, row_ptr, col, val, label)
    _check_all(ds, ref, oracle.LOGISTIC, rng, draws=1)
    ds.close()


def test_empty_dataset(ctx):
    ds, ref = _pair(ctx, 0, 4, np.zeros(1, np.int64), np.zeros(0, np.int32), np.zeros(0), np.zeros(0))
    f, g = ds.value_and_gradient(oracle.LOGISTIC, np.ones(4))
    assert f == 0.0 and np.array_equal(g, np.zeros(4))
    assert np.array_equal(ds.hessian_vector(oracle.LOGISTIC, np.ones(4), np.ones(4)), np.zeros(4))
    ds.close()


def test_offsets_weights_labels_can_be_replaced(ctx):
    # Coordinate.updateModel re-offsets the same packed rows every pass (Coordinate.scala:59-62)
    rng = np.random.default_rng(40)
    n, d = 3000, 200
    row_ptr, col, val = util.random_csr(rng, n, d, 12)
    label = util.labels_for(rng, 0, n)
    ds, _ = _pair(ctx, n, d, row_ptr, col, val, label)
    w = rng.normal(scale=0.3, size=d)
    v = rng.normal(size=d)
    hv_before = ds.hessian_vector(oracle.LOGISTIC, w, v, tag=5)
    for _ in range(2):
        off, wt = rng.normal(size=n), rng.uniform(0.5, 2, size=n)
        lab = util.labels_for(rng, 0, n)
        ds.set_offsets(off)
        ds.set_weights(wt)
        ds.set_labels(lab)
        ref = oracle.Data(n, d, row_ptr, col, val, lab, off, wt)
        f_ref, g_ref = oracle.value_and_gradient(ref, oracle.LOGISTIC, w)
        f, g = ds.value_and_gradient(oracle.LOGISTIC, w)
        assert abs(f - f_ref) <= TOL * abs(f_ref) and util.rel_err(g, g_ref) <= TOL
        # same tag, but the rows changed: the cache must have been dropped
        hv = ds.hessian_vector(oracle.LOGISTIC, w, v, tag=5)
        assert util.rel_err(hv, oracle.hessian_vector(ref, oracle.LOGISTIC, w, v)) <= TOL
    ds.set_offsets(None)
    ds.set_weights(None)
    ds.set_labels(label)
    assert util.rel_err(ds.hessian_vector(oracle.LOGISTIC, w, v, tag=5), hv_before) <= TOL
    ds.close()


def test_scores(ctx):
    rng = np.random.default_rng(50)
    n, d = 5000, 300
    row_ptr, col, val = util.random_csr(rng, n, d, 20)
    ds, _ = _pair(ctx, n, d, row_ptr, col, val, np.zeros(n), rng.normal(size=n))
    w = rng.normal(size=d)
    X = util.csr_to_dense(n, d, row_ptr, col, val)
    assert util.rel_err(ds.scores(w), X @ w) <= TOL  # no offset: Coefficients.computeScore
    ds.close()


def test_requires_become_illegal_argument(ctx):
    import photon_ml_b200 as pml
    rng = np.random.default_rng(60)
    n, d = 50, 6
    row_ptr, col, val = util.random_csr(rng, n, d, 3)
    ds, _ = _pair(ctx, n, d, row_ptr, col, val, np.zeros(n))
    with pytest.raises(pml.IllegalArgumentError):  # coefficient size (ValueAndGradientAggregator.scala:141-143)
        ds.value_and_gradient(oracle.LOGISTIC, np.zeros(d + 1))
    with pytest.raises(pml.IllegalArgumentError):  # multiply vector size (HessianVectorAggregator.scala:63)
        ds.hessian_vector(oracle.LOGISTIC, np.zeros(d), np.zeros(d - 1))
    with pytest.raises(pml.IllegalArgumentError):  # factors size (ValueAndGradientAggregator.scala:100)
        ds.set_normalization(factors=np.ones(d + 2), intercept=0)
    with pytest.raises(pml.IllegalArgumentError):  # intercept factor (:96-99)
        ds.set_normalization(factors=np.full(d, 2.0), intercept=0)
    with pytest.raises(pml.IllegalArgumentError):  # intercept shift (:109-112)
        ds.set_normalization(shifts=np.ones(d), intercept=0)
    with pytest.raises(pml.IllegalArgumentError):  # shift without intercept (NormalizationContext.scala:44)
        ds.set_normalization(shifts=np.zeros(d))
    with pytest.raises(pml.IllegalArgumentError):  # hinge is not a TwiceDiffFunction
        ds.hessian_vector(oracle.SMOOTHED_HINGE, np.zeros(d), np.zeros(d))
    with pytest.raises(pml.IllegalArgumentError):
        ds.hessian_diagonal(oracle.SMOOTHED_HINGE, np.zeros(d))
    with pytest.raises(pml.IllegalArgumentError):
        ds.value_and_gradient(9, np.zeros(d))
    bad_col = col.copy()
    bad_col[0] = d
    with pytest.raises(pml.IllegalArgumentError):  # features.size == dim
        pml.Dataset.from_csr(ctx, n, d, row_ptr, bad_col, val, np.zeros(n))
    bad_ptr = row_ptr.copy()
    bad_ptr[3] = bad_ptr[4] + 1
    with pytest.raises(pml.IllegalArgumentError):
        pml.Dataset.from_csr(ctx, n, d, bad_ptr, col, val, np.zeros(n))
    # the handle is still usable after refused calls
    f, _ = ds.value_and_gradient(oracle.LOGISTIC, np.zeros(d))
    assert abs(f - n * np.log(2.0)) <= 1e-12 * n
    ds.close()


def test_nan_and_inf_propagate(ctx):
    # Poisson has no overflow guard (PoissonLossFunction.scala:41); NaN features are not an error
    rng = np.random.default_rng(70)
    n, d = 100, 5
    row_ptr, col, val = util.random_csr(rng, n, d, 3)
    ds, ref = _pair(ctx, n, d, row_ptr, col, val, np.ones(n))
    f, g = ds.value_and_gradient(oracle.POISSON, np.full(d, 1e4))
    f_ref, g_ref = oracle.value_and_gradient(ref, oracle.POISSON, np.full(d, 1e4))
    assert np.isinf(f) or np.isnan(f)
    assert (np.isinf(f_ref) or np.isnan(f_ref))
    val2 = val.copy()
    val2[0] = np.nan
    ds2, _ = _pair(ctx, n, d, row_ptr, col, val2, np.ones(n))
    f2, _ = ds2.value_and_gradient(oracle.LOGISTIC, np.ones(d))
    assert np.isnan(f2)
    ds.close()
    ds2.close()


def test_bitwise_reproducible(ctx):
    rng = np.random.default_rng(80)
    n, d = 8000, 900
    row_ptr, col, val = util.random_csr(rng, n, d, 15, skew=1.2)
    ds, _ = _pair(ctx, n, d, row_ptr, col, val, util.labels_for(rng, 0, n))
    w = rng.normal(size=d)
    a = ds.value_and_gradient(oracle.LOGISTIC, w)
    for _ in range(3):
        b = ds.value_and_gradient(oracle.LOGISTIC, w)
        assert a[0] == b[0] and np.array_equal(a[1], b[1])  # no atomics: fixed reduction shapes
    ds.close()


def test_device_buffer_entry_points_and_rank_split(ctx):
    # the multi-GPU split on one device: two row shards -> raw partials -> sum -> finalize once
    import torch
    import photon_ml_b200 as pml
    rng = np.random.default_rng(90)
    n, d = 7000, 400
    row_ptr, col, val = util.random_csr(rng, n, d, 11, skew=1.1)
    label = util.labels_for(rng, 0, n)
    off = rng.normal(scale=0.1, size=n)
    f_, s_, icpt = util.normalization_context("STANDARDIZATION", n, d, row_ptr, col, val, d - 1)
    ref = oracle.Data(n, d, row_ptr, col, val, label, off)
    cut = 3000
    shards = []
    for lo, hi in ((0, cut), (cut, n)):
        rp = row_ptr[lo:hi + 1] - row_ptr[lo]
        sl = slice(row_ptr[lo], row_ptr[hi])
        ds = pml.Dataset.from_csr(ctx, hi - lo, d, rp, col[sl], val[sl], label[lo:hi], off[lo:hi])
        ds.set_normalization(f_, s_, icpt)
        shards.append(ds)
    dev = torch.device("cuda", 0)
    w = torch.tensor(rng.normal(scale=0.2, size=d), device=dev)
    v = torch.tensor(rng.normal(size=d), device=dev)
    raws = [torch.empty(d + 2, dtype=torch.float64, device=dev) for _ in shards]
    torch.cuda.synchronize()
    for ds, raw in zip(shards, raws):
        ds.value_and_gradient_raw_dev(oracle.LOGISTIC, w.data_ptr(), raw.data_ptr())
    ctx.synchronize()
    total = raws[0] + raws[1]
    out_v = torch.empty(1, dtype=torch.float64, device=dev)
    out_g = torch.empty(d, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    shards[0].finalize_dev(oracle.LOGISTIC, total.data_ptr(), out_v.data_ptr(), out_g.data_ptr())
    ctx.synchronize()
    f_ref, g_ref = oracle.value_and_gradient(ref, oracle.LOGISTIC, w.cpu().numpy(), f_, s_, icpt)
    assert abs(out_v.item() - f_ref) <= TOL * abs(f_ref)
    assert util.rel_err(out_g.cpu().numpy(), g_ref) <= TOL
    for ds, raw in zip(shards, raws):
        ds.hessian_vector_raw_dev(oracle.LOGISTIC, w.data_ptr(), v.data_ptr(), raw.data_ptr())
    ctx.synchronize()
    total = raws[0] + raws[1]
    torch.cuda.synchronize()
    shards[0].finalize_dev(oracle.LOGISTIC, total.data_ptr(), None, out_g.data_ptr())
    ctx.synchronize()
    hv_ref = oracle.hessian_vector(ref, oracle.LOGISTIC, w.cpu().numpy(), v.cpu().numpy(), f_, s_, icpt)
    assert util.rel_err(out_g.cpu().numpy(), hv_ref) <= TOL
    # single-shard convenience calls
    one = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label, off)
    one.value_and_gradient_dev(oracle.LOGISTIC, w.data_ptr(), out_v.data_ptr(), out_g.data_ptr())
    ctx.synchronize()
    f_ref, g_ref = oracle.value_and_gradient(ref, oracle.LOGISTIC, w.cpu().numpy())
    assert abs(out_v.item() - f_ref) <= TOL * abs(f_ref) and util.rel_err(out_g.cpu().numpy(), g_ref) <= TOL
    one.hessian_vector_dev(oracle.LOGISTIC, w.data_ptr(), v.data_ptr(), out_g.data_ptr())
    ctx.synchronize()
    assert util.rel_err(out_g.cpu().numpy(), oracle.hessian_vector(ref, oracle.LOGISTIC, w.cpu().numpy(), v.cpu().numpy())) <= TOL
    for ds in shards + [one]:
        ds.close()


def test_medium_problem_properties_and_oracle(ctx):
    # ~2M entries: still seconds for the oracle; also checks Hv linearity and Hv vs finite differences
    rng = np.random.default_rng(100)
    n, d, k = 60000, 20000, 35
    row_ptr, col, val = util.fast_random_csr(rng, n, d, k, skew=1.1)
    z = rng.normal(scale=0.5, size=n)
    label = util.labels_for(rng, 1, n, z)
    ds, ref = _pair(ctx, n, d, row_ptr, col, val, label)
    assert ds.info().value_bytes == 4
    w = rng.normal(scale=0.02, size=d)
    v1, v2 = rng.normal(size=d), rng.normal(size=d)
    f_ref, g_ref = oracle.value_and_gradient(ref, oracle.POISSON, w, parts=4)
    f, g = ds.value_and_gradient(oracle.POISSON, w)
    assert abs(f - f_ref) <= TOL * abs(f_ref) and util.rel_err(g, g_ref) <= TOL
    h1 = ds.hessian_vector(oracle.POISSON, w, v1, tag=1)
    h2 = ds.hessian_vector(oracle.POISSON, w, v2, tag=1)
    h12 = ds.hessian_vector(oracle.POISSON, w, 2.0 * v1 - 3.0 * v2, tag=1)
    assert util.rel_err(h12, 2.0 * h1 - 3.0 * h2) <= 1e-12
    assert util.rel_err(h1, oracle.hessian_vector(ref, oracle.POISSON, w, v1)) <= TOL
    eps = 1e-6
    gp = ds.value_and_gradient(oracle.POISSON, w + eps * v1)[1]
    gm = ds.value_and_gradient(oracle.POISSON, w - eps * v1)[1]
    assert util.rel_err((gp - gm) / (2 * eps), h1) <= 1e-5
    ds.close()


def test_two_contexts_from_two_threads():
    # Handles on different contexts may be used concurrently from different threads (TestNG parallel=methods,
    # executor task threads); each context has its own stream and lock.
    import threading
    import photon_ml_b200 as pml
    rng = np.random.default_rng(200)
    n, d = 4000, 150
    problems = []
    for _ in range(2):
        row_ptr, col, val = util.random_csr(rng, n, d, 10, skew=1.1)
        label = util.labels_for(rng, 0, n)
        w = rng.normal(scale=0.3, size=d)
        ref = oracle.Data(n, d, row_ptr, col, val, label)
        problems.append((row_ptr, col, val, label, w, oracle.value_and_gradient(ref, oracle.LOGISTIC, w)))
    errors = []

    def work(k):
        try:
            row_ptr, col, val, label, w, (f_ref, g_ref) = problems[k]
            c = pml.Context(0)
            ds = pml.Dataset.from_csr(c, n, d, row_ptr, col, val, label)
            for _ in range(50):
                f, g = ds.value_and_gradient(oracle.LOGISTIC, w)
                assert abs(f - f_ref) <= TOL * abs(f_ref) and util.rel_err(g, g_ref) <= TOL
            ds.close()
            c.close()
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors


@pytest.mark.parametrize("loss", LOSSES + [oracle.SMOOTHED_HINGE])
def test_training_loss_evaluator_on_device_scores(ctx, loss):
    """pml_loss_from_scores_dev: sum wt * l(score + offset, y) for scores that are already on the device equals the
    objective value at the coefficients the scores came from (Evaluator.evaluate, lib/evaluation/Evaluator.scala:35-45)."""
    import torch
    rng = np.random.default_rng(60 + loss)
    n, d = 2500, 300
    row_ptr, col, val = util.random_csr(rng, n, d, 7, skew=1.1, empty_rows=0.05)
    label = util.labels_for(rng, loss, n)
    offset, weight = rng.normal(scale=0.3, size=n), rng.uniform(0.5, 2.0, size=n)
    ds, ref = _pair(ctx, n, d, row_ptr, col, val, label, offset, weight)
    w = rng.normal(scale=0.2, size=d)
    f_ref, _ = oracle.value_and_gradient(ref, loss, w)
    dev = torch.device("cuda", 0)
    with torch.cuda.stream(torch.cuda.ExternalStream(ctx.stream, device=dev)):
        t_w = torch.from_numpy(w).to(dev)
        t_scores = torch.empty(n, dtype=torch.float64, device=dev)
        t_value = torch.zeros(1, dtype=torch.float64, device=dev)
        ds.scores_dev(t_w.data_ptr(), t_scores.data_ptr())
        ds.loss_from_scores_dev(loss, t_scores.data_ptr(), t_value.data_ptr())
        ctx.synchronize()
        assert abs(float(t_value.cpu()) - f_ref) <= TOL * abs(f_ref)
        # and the evaluation that follows is not disturbed by the evaluator's use of the scratch
        f, g = ds.value_and_gradient(loss, w)
        f_ref, g_ref = oracle.value_and_gradient(ref, loss, w)
        assert abs(f - f_ref) <= TOL * abs(f_ref) and util.rel_err(g, g_ref) <= TOL
    ds.close()
