"""The reference's golden vectors and known-answer tests run on the CUDA path itself (the oracle is pinned to them
in test_oracle_golden.py; here the device code is).  Fixtures: tests/golden/reference_kats.json, heart.txt.  `-m gpu`."""
import math
import os

import numpy as np
import pytest

import oracle
from tests import util

pytestmark = pytest.mark.gpu

K = util.kats()
LOSSES = [oracle.LOGISTIC, oracle.POISSON, oracle.SQUARED]


@pytest.fixture(scope="module")
def ctx():
    import photon_ml_b200 as pml
    c = pml.Context(0)
    yield c
    c.close()


def _one_row(ctx, features, label, offset=0.0):
    """A data set of one LabeledPoint with a dense feature vector."""
    import photon_ml_b200 as pml
    d = len(features)
    return pml.Dataset.from_csr(ctx, 1, d, np.array([0, d]), np.arange(d, dtype=np.int32), np.asarray(features, np.float64),
                                np.array([label]), np.array([offset]))


def test_log1p_exp_kats_on_the_device(ctx):
    # MathUtilsTest.scala:29-34: with label 0 the logistic loss of one row IS log1pExp(margin)
    for x, want in K["log1pExp"]["cases"]:
        ds = _one_row(ctx, [1.0], 0.0)
        value, _ = ds.value_and_gradient(oracle.LOGISTIC, np.array([float(x)]))
        assert abs(value - want) <= K["log1pExp"]["tolerance"], (x, value, want)
        ds.close()


def test_logistic_identities_on_the_device(ctx):
    # LogisticLossFunctionTest.scala:29-93: l, l', l'' against the naive formulas at the reference's margins
    tol = K["logistic_identities"]["tolerance"]
    c = K["logistic_identities"]["testCalculate"]
    m = c["offset"] + float(np.dot(c["features"], c["coefficients"]))
    for label, want in ((1.0, math.log(1 + math.exp(-m))), (0.0, math.log(1 + math.exp(m)))):
        ds = _one_row(ctx, c["features"], label, c["offset"])
        value, _ = ds.value_and_gradient(oracle.LOGISTIC, np.asarray(c["coefficients"], np.float64))
        assert abs(value - want) <= tol * max(1.0, abs(want))
        ds.close()
    g = K["logistic_identities"]["testGradient"]
    m = g["offset"] + float(np.dot(g["features"], g["coefficients"]))
    x = np.asarray(g["features"])
    for label, dl in ((1.0, -1.0 / (1.0 + math.exp(m))), (0.0, 1.0 / (1.0 + math.exp(-m)))):
        ds = _one_row(ctx, g["features"], label, g["offset"])
        _, grad = ds.value_and_gradient(oracle.LOGISTIC, np.asarray(g["coefficients"], np.float64))
        assert np.max(np.abs(grad - dl * x)) <= tol * max(1.0, np.max(np.abs(dl * x)))
        ds.close()
    for h in K["logistic_identities"]["testHessianVector"]:
        m = float(np.dot(h["features"], h["coefficients"]))
        s = 1.0 / (1.0 + math.exp(-m))
        x = np.asarray(h["features"], np.float64)
        ds = _one_row(ctx, h["features"], 1.0)
        diag = ds.hessian_diagonal(oracle.LOGISTIC, np.asarray(h["coefficients"], np.float64))
        assert np.max(np.abs(diag - s * (1 - s) * x * x)) <= tol * max(1.0, np.max(np.abs(x * x)))
        v = np.ones(len(x))
        hv = ds.hessian_vector(oracle.LOGISTIC, np.asarray(h["coefficients"], np.float64), v)
        assert np.max(np.abs(hv - s * (1 - s) * float(x.dot(v)) * x)) <= tol * max(1.0, np.max(np.abs(x)) * abs(float(x.dot(v))))
        ds.close()


def _trivial(ctx):
    import photon_ml_b200 as pml
    rows = np.asarray(K["trivial_labeled_points"]["rows"])
    X = np.hstack([rows[:, 1:], np.ones((len(rows), 1))])
    row_ptr, col, val = util.dense_to_csr(X, keep_zeros=True)   # DenseVector rows: zeros are stored entries
    return pml.Dataset.from_csr(ctx, len(rows), 3, row_ptr, col, val, rows[:, 0]), (len(rows), row_ptr, col, val)


@pytest.mark.parametrize("optimizer", ["LBFGS", "TRON"])
def test_ridge_golden_from_scikit_learn_trained_on_the_device(ctx, optimizer):
    # GameEstimatorIntegTest.scala:61-105: squared loss, L2 0.3 (intercept penalised too), tolerance 1e-11
    from photon_ml_b200.function import DistributedGLMLossFunction, NoNormalization
    from photon_ml_b200.optimization import LBFGS, TRON
    ds, _ = _trivial(ctx)
    opt = (LBFGS(NoNormalization(), tolerance=1e-11, maxNumIterations=200) if optimizer == "LBFGS"
           else TRON(NoNormalization(), tolerance=1e-11, maxNumIterations=50))
    w, _ = opt.optimize(DistributedGLMLossFunction.create(oracle.SQUARED, l2RegWeight=K["ridge_golden"]["l2"]), np.zeros(3), ds)
    # 5e-10: the typed-in scikit-learn numbers are themselves 1.6e-10 from the exact optimum (test_oracle_golden.py)
    assert np.max(np.abs(w - K["ridge_golden"]["coefficients"])) <= 5e-10, w
    ds.close()


@pytest.mark.parametrize("kind", K["normalization_golden"]["normalization_types"])
def test_normalization_golden_trained_on_the_device(ctx, kind):
    # GameEstimatorIntegTest.scala:129-179: lambda = 0, every normalization type -> the same model in the original space
    from photon_ml_b200.function import DistributedGLMLossFunction, NormalizationContext
    from photon_ml_b200.optimization import TRON
    ds, (n, row_ptr, col, val) = _trivial(ctx)
    f, s, icpt = util.normalization_context(kind, n, 3, row_ptr, col, val, 2)
    norm = NormalizationContext(f, s, icpt)
    opt = TRON(norm, tolerance=1e-12, maxNumIterations=50)
    wt, _ = opt.optimize(DistributedGLMLossFunction.create(oracle.SQUARED), np.zeros(3), ds)
    w = norm.modelToOriginalSpace(wt)
    assert np.max(np.abs(w - K["normalization_golden"]["coefficients"])) <= K["normalization_golden"]["tolerance"], w
    ds.close()


@pytest.mark.parametrize("loss", LOSSES)
@pytest.mark.parametrize("optimizer", ["LBFGS", "TRON"])
def test_heart_training_with_a_context_equals_training_on_standardised_data(ctx, loss, optimizer):
    # NormalizationContextIntegTest.scala:165-262: heart.txt under a STANDARDIZATION context vs the explicitly
    # standardised features without a context; same optimizer, tolerance 1e-6, <= 100 iterations.  (L2 = 1 is this
    # test's addition: it keeps both runs well-conditioned so that they stop at the same iteration.)
    import photon_ml_b200 as pml
    from photon_ml_b200.function import DistributedGLMLossFunction, NoNormalization, NormalizationContext
    from photon_ml_b200.optimization import LBFGS, TRON
    n, d, row_ptr, col, val, label = util.read_libsvm(os.path.join(util.GOLDEN, "heart.txt"), d=13, add_intercept=True)
    lab = label if loss != oracle.POISSON else label * 3.0
    f, s, icpt = util.normalization_context("STANDARDIZATION", n, d, row_ptr, col, val, d - 1)
    X = util.csr_to_dense(n, d, row_ptr, col, val)
    rp2, c2, v2 = util.dense_to_csr((X - s) * f, keep_zeros=True)
    raw = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, lab)
    explicit = pml.Dataset.from_csr(ctx, n, d, rp2, c2, v2, lab)

    def make(norm):
        return (LBFGS(norm, tolerance=1e-6, maxNumIterations=100) if optimizer == "LBFGS"
                else TRON(norm, tolerance=1e-6, maxNumIterations=100))
    obj = DistributedGLMLossFunction.create(loss, l2RegWeight=1.0)
    a, b = make(NormalizationContext(f, s, icpt)), make(NoNormalization())
    wa, fa = a.optimize(obj, np.zeros(d), raw)
    wb, fb = b.optimize(DistributedGLMLossFunction.create(loss, l2RegWeight=1.0), np.zeros(d), explicit)
    assert a.currentState.iter == b.currentState.iter
    assert abs(fa - fb) <= 1e-9 * abs(fb)
    assert np.max(np.abs(wa - wb)) <= 1e-6 * max(1.0, np.max(np.abs(wb)))
    raw.close()
    explicit.close()
