"""Optimizer drivers (photon_ml_b200.optimization) on the CPU: the reference's own optimizer checks
(lib-test/optimization/LBFGSTest.scala:50-67, OWLQNTest.scala:57-79, lib-integ/optimization/
OptimizerIntegTest.scala:70-117) plus the converged-coefficient goldens driven through the oracle."""
import os

import numpy as np
import pytest

import oracle
from photon_ml_b200.function import NoNormalization, NormalizationContext
from photon_ml_b200.optimization import (FUNCTION_VALUES_CONVERGED, GRADIENT_CONVERGED, LBFGS, OWLQN, TRON)
from tests import util
from tests.objectives import OracleGLMObjective, QuadraticObjective

K = util.kats()


@pytest.mark.parametrize("make", [lambda: LBFGS(NoNormalization(), tolerance=1e-9),
                                  lambda: TRON(NoNormalization(), tolerance=1e-9, maxNumIterations=50)])
def test_quadratic_reaches_centroid(make):
    opt = make()
    rng = np.random.default_rng(0)
    for _ in range(5):
        x0 = rng.uniform(-10, 10, size=10)
        x, f = opt.optimize(QuadraticObjective(), x0, None)
        assert np.max(np.abs(x - 4.0)) <= 1e-6 and f <= 1e-10
        assert opt.convergenceReason in (FUNCTION_VALUES_CONVERGED, GRADIENT_CONVERGED)
        losses = [s[1] for s in opt.trackedStates]
        assert all(b <= a + 1e-12 for a, b in zip(losses, losses[1:]))   # monotone decrease


@pytest.mark.parametrize("l1", [0.1, 1.0, 4.0, 10.0])
def test_owlqn_shrinkage(l1):
    # minimise sum (x-4)^2 + l1*|x|: x* = max(4 - l1/2, 0)  (OWLQNTest.scala:44-55)
    opt = OWLQN(l1, NoNormalization(), tolerance=1e-10)
    x, _ = opt.optimize(QuadraticObjective(), np.zeros(6), None)
    assert np.max(np.abs(x - max(4.0 - l1 / 2.0, 0.0))) <= 1e-5


def test_box_constraints_are_projected():
    opt = LBFGS(NoNormalization(), constraintMap={0: (0.0, 1.0), 3: (5.0, 6.0)})
    x, _ = opt.optimize(QuadraticObjective(), np.zeros(5), None)
    assert x[0] == 1.0 and x[3] == 5.0 and abs(x[1] - 4.0) <= 1e-6


def _trivial():
    rows = np.asarray(K["trivial_labeled_points"]["rows"])
    X = np.hstack([rows[:, 1:], np.ones((len(rows), 1))])
    row_ptr, col, val = util.dense_to_csr(X, keep_zeros=True)
    return oracle.Data(len(rows), 3, row_ptr, col, val, rows[:, 0]), (row_ptr, col, val)


@pytest.mark.parametrize("kind", ["LBFGS", "TRON"])
def test_ridge_golden_converged_coefficients(kind):
    # GameEstimatorIntegTest.scala:61-105 (LBFGS there); tolerance: see tests/test_oracle_golden.py
    data, _ = _trivial()
    obj = OracleGLMObjective(oracle.SQUARED, l2RegWeight=0.3)
    opt = LBFGS(NoNormalization(), tolerance=1e-11) if kind == "LBFGS" else TRON(NoNormalization(), tolerance=1e-11, maxNumIterations=100)
    w, _ = opt.optimize(obj, np.zeros(3), data)
    assert np.max(np.abs(w - K["ridge_golden"]["coefficients"])) <= 5e-10


@pytest.mark.parametrize("norm_kind", K["normalization_golden"]["normalization_types"])
def test_normalization_golden_converged_coefficients(norm_kind):
    # GameEstimatorIntegTest.scala:129-179: train in the transformed space, map back, same optimum
    data, (row_ptr, col, val) = _trivial()
    f, s, icpt = util.normalization_context(norm_kind, data.n, 3, row_ptr, col, val, 2)
    ctx = NormalizationContext(f, s, icpt)
    opt = LBFGS(ctx, tolerance=1e-11)
    wt, _ = opt.optimize(OracleGLMObjective(oracle.SQUARED), np.zeros(3), data)
    w = ctx.modelToOriginalSpace(wt)
    assert np.max(np.abs(w - K["normalization_golden"]["coefficients"])) <= K["normalization_golden"]["tolerance"]


@pytest.mark.parametrize("loss", [oracle.LOGISTIC, oracle.SQUARED, oracle.POISSON])
@pytest.mark.parametrize("kind", ["LBFGS", "TRON"])
def test_heart_training_with_context_equals_training_on_transformed_data(loss, kind):
    # NormalizationContextIntegTest.scala:165-262
    n, d, row_ptr, col, val, label = util.read_libsvm(os.path.join(util.GOLDEN, "heart.txt"), d=13, add_intercept=True)
    f, s, icpt = util.normalization_context("STANDARDIZATION", n, d, row_ptr, col, val, d - 1)
    X = util.csr_to_dense(n, d, row_ptr, col, val)
    rp2, c2, v2 = util.dense_to_csr((X - s) * f, keep_zeros=True)
    lab = label if loss != oracle.POISSON else label * 2.0
    raw, explicit = oracle.Data(n, d, row_ptr, col, val, lab), oracle.Data(n, d, rp2, c2, v2, lab)
    ctx = NormalizationContext(f, s, icpt)

    def make(c):
        return LBFGS(c, tolerance=1e-6) if kind == "LBFGS" else TRON(c, tolerance=1e-6, maxNumIterations=100)

    a, b = make(ctx), make(NoNormalization())
    wa, fa = a.optimize(OracleGLMObjective(loss), np.zeros(d), raw)
    wb, fb = b.optimize(OracleGLMObjective(loss), np.zeros(d), explicit)
    assert abs(fa - fb) <= 1e-9 * abs(fb)
    assert np.max(np.abs(wa - wb)) <= 1e-7 * max(1.0, np.max(np.abs(wb)))
    assert a.currentState.iter == b.currentState.iter


@pytest.mark.parametrize("l1", [0.0, 2.0])
def test_lbfgs_and_owlqn_take_numpy_or_torch_vectors(l1):
    """The drivers keep their state in whatever vector type the objective speaks (numpy on the host-buffer ABI, torch
    tensors for device-resident state): same iterates either way."""
    import torch
    from photon_ml_b200.function import NoNormalization
    from photon_ml_b200.optimization import LBFGS, OWLQN
    rng = np.random.default_rng(0)
    A, b = rng.normal(size=(40, 15)), rng.normal(size=40)

    class LeastSquares:
        twice_differentiable = True

        def calculate(self, data, x, ctx, want_gradient=True):
            if isinstance(x, np.ndarray):
                r = A @ x - b
                return float(0.5 * r.dot(r)), A.T @ r
            r = torch.from_numpy(A) @ x - torch.from_numpy(b)
            return float(0.5 * r.dot(r)), torch.from_numpy(A).T @ r

    def make():
        return OWLQN(l1, NoNormalization()) if l1 else LBFGS(NoNormalization())
    host, dev = make(), make()
    w_host, f_host = host.optimize(LeastSquares(), np.zeros(15), None)
    w_dev, f_dev = dev.optimize(LeastSquares(), torch.zeros(15, dtype=torch.float64), None)
    assert host.currentState.iter == dev.currentState.iter and host.evaluations == dev.evaluations
    assert abs(f_host - f_dev) <= 1e-12 * abs(f_host)
    assert np.max(np.abs(w_host - w_dev.numpy())) <= 1e-12
    if l1:
        assert int((w_host == 0).sum()) == int((w_dev.numpy() == 0).sum()) > 0
