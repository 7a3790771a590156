"""The reference's own objective-function tests, restated against the CUDA objective through the mirror of its classes
(photon_ml_b200.function): SingleNodeObjectiveFunctionTest.scala:416-512 and DistributedObjectiveFunctionTest.scala:
449-573 check value / gradient / Hessian-vector consistency by central differences (delta 1e-6, tolerance 1e-3) on
generated "benign" and "outlier" data, with and without weights, with and without L2.  `-m gpu`."""
import numpy as np
import pytest

import oracle
from tests import util

pytestmark = pytest.mark.gpu

LOSSES = [oracle.LOGISTIC, oracle.POISSON, oracle.SQUARED]
DELTA, TOL = 1e-6, 1e-3


def _generated(rng, loss, outlier, weighted, n=25, d=5):
    # the shapes of SingleNodeObjectiveFunctionTest.scala:516-530: dimension 5, 25 rows
    if outlier:
        X = np.where(rng.random((n, d)) < 0.9, rng.normal(scale=1e-3, size=(n, d)), rng.choice([-1.0, 1.0], (n, d)))
    else:
        X = rng.uniform(-1, 1, size=(n, d))
    row_ptr, col, val = util.dense_to_csr(X, keep_zeros=True)
    label = util.labels_for(rng, loss, n)
    weight = rng.uniform(0.1, 10.0, size=n) if weighted else None
    offset = rng.normal(scale=0.1, size=n)
    return n, d, row_ptr, col, val, label, offset, weight


def _objective(loss, l2):
    from photon_ml_b200.function import DistributedGLMLossFunction, DistributedSmoothedHingeLossFunction
    if loss == oracle.SMOOTHED_HINGE:
        return DistributedSmoothedHingeLossFunction.create(l2RegWeight=l2 or None)
    return DistributedGLMLossFunction.create(loss, l2RegWeight=l2 or None)


@pytest.mark.parametrize("loss", LOSSES + [oracle.SMOOTHED_HINGE])
@pytest.mark.parametrize("outlier", [False, True])
@pytest.mark.parametrize("weighted", [False, True])
def test_gradient_and_hessian_vector_by_central_differences(loss, outlier, weighted):
    import photon_ml_b200 as pml
    from photon_ml_b200.function import NoNormalization
    rng = np.random.default_rng(2000 * loss + 10 * outlier + weighted)
    n, d, row_ptr, col, val, label, offset, weight = _generated(rng, loss, outlier, weighted)
    ctx = pml.Context(0)
    # the generated values are arbitrary doubles: stored as fp64, exactly what the reference holds
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label, offset, weight)
    norm = NoNormalization()
    for l2 in (0.0, 100.0):
        obj = _objective(loss, l2)
        for _ in range(5):
            w = rng.normal(size=d)
            f0, g = obj.calculate(ds, w, norm)
            assert f0 == obj.value(ds, w, norm) and np.array_equal(g, obj.gradient(ds, w, norm))
            for j in range(d):
                e = np.zeros(d)
                e[j] = DELTA
                num = (obj.value(ds, w + e, norm) - obj.value(ds, w - e, norm)) / (2 * DELTA)
                assert abs(num - g[j]) <= TOL * max(1.0, abs(g[j])), (l2, j, num, g[j])
            if loss == oracle.SMOOTHED_HINGE:
                with pytest.raises(pml.IllegalArgumentError):
                    obj.hessianVector(ds, w, w, norm)          # DiffFunction only (TRON's cast fails in the reference)
                continue
            v = rng.normal(size=d)
            hv = obj.hessianVector(ds, w, v, norm)
            num = (obj.gradient(ds, w + DELTA * v, norm) - obj.gradient(ds, w - DELTA * v, norm)) / (2 * DELTA)
            assert np.all(np.abs(num - hv) <= TOL * np.maximum(1.0, np.abs(hv)))
            # the diagonal is e_j' H e_j
            diag = obj.hessianDiagonal(ds, w)
            for j in range(d):
                e = np.zeros(d)
                e[j] = 1.0
                assert abs(obj.hessianVector(ds, w, e, norm)[j] - diag[j]) <= 1e-9 * max(1.0, abs(diag[j]))
    ds.close()
    ctx.close()
