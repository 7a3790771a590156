"""Convergence parity: the same optimizer driver run on the CUDA objective (through the C ABI) and on
the CPU oracle objective must take the same number of iterations and land on the same coefficients
(north star: <= 1e-6 on converged coefficients, identical iteration counts on well-conditioned
problems).  `-m gpu`."""
import os

import numpy as np
import pytest

import oracle
from tests import util
from tests.objectives import OracleGLMObjective

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import photon_ml_b200 as pml
    c = pml.Context(0)
    yield c
    c.close()


def _a1a():
    return util.read_libsvm(os.path.join(util.GOLDEN, "a9a_1605.libsvm"), d=123, add_intercept=True)


@pytest.mark.parametrize("lam", [100.0, 10.0, 1.0, 0.1])
def test_config1_a1a_logistic_l2_lbfgs(ctx, lam):
    # BASELINE.json configs[0] / README.md:238-254: a1a logistic, L2, LBFGS m=10 tol=1e-7, <= 50 iterations
    import photon_ml_b200 as pml
    from photon_ml_b200.function import DistributedGLMLossFunction, NoNormalization
    from photon_ml_b200.optimization import LBFGS
    n, d, row_ptr, col, val, label = _a1a()
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
    ref = oracle.Data(n, d, row_ptr, col, val, label)
    gpu_opt = LBFGS(NoNormalization(), maxNumIterations=50)
    cpu_opt = LBFGS(NoNormalization(), maxNumIterations=50)
    w_gpu, f_gpu = gpu_opt.optimize(DistributedGLMLossFunction.create(pml.LOGISTIC, l2RegWeight=lam), np.zeros(d), ds)
    w_cpu, f_cpu = cpu_opt.optimize(OracleGLMObjective(oracle.LOGISTIC, l2RegWeight=lam), np.zeros(d), ref)
    assert gpu_opt.currentState.iter == cpu_opt.currentState.iter
    assert gpu_opt.evaluations == cpu_opt.evaluations
    assert gpu_opt.convergenceReason == cpu_opt.convergenceReason
    assert abs(f_gpu - f_cpu) <= 1e-9 * abs(f_cpu)
    assert np.max(np.abs(w_gpu - w_cpu)) <= 1e-6 * max(1.0, np.max(np.abs(w_cpu)))
    ds.close()


@pytest.mark.parametrize("loss,l2", [(oracle.POISSON, 100.0), (oracle.LOGISTIC, 100.0)])
def test_tron_hessian_vector_path(ctx, loss, l2):
    # configs[1] in miniature: TRON (tol 1e-5, <= 15 iterations, <= 20 CG each) on a skewed sparse problem.
    # "Well-conditioned" matters: with l2 = 1 the Poisson problem is still descending when TRON hits its
    # 15-iteration cap, and the trust-region dynamics amplify 1e-16 evaluation differences about tenfold per
    # iteration (same iteration and CG counts, objective values 1e-4 apart at the cap); l2 = 100 converges.
    import photon_ml_b200 as pml
    from photon_ml_b200.function import DistributedGLMLossFunction, NoNormalization
    from photon_ml_b200.optimization import TRON
    rng = np.random.default_rng(21)
    n, d, k = 20000, 3000, 20
    row_ptr, col, val = util.fast_random_csr(rng, n, d, k, skew=1.1)
    w_true = rng.normal(scale=0.3, size=d) * (rng.random(d) < 0.05)
    z = np.array([val[row_ptr[i]:row_ptr[i + 1]].dot(w_true[col[row_ptr[i]:row_ptr[i + 1]]]) for i in range(n)]).clip(-5, 5)
    label = util.labels_for(rng, loss, n, z) if loss == oracle.POISSON else (rng.random(n) < 1 / (1 + np.exp(-z))).astype(float)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
    ref = oracle.Data(n, d, row_ptr, col, val, label)
    gpu_opt, cpu_opt = TRON(NoNormalization()), TRON(NoNormalization())
    w_gpu, f_gpu = gpu_opt.optimize(DistributedGLMLossFunction.create(loss, l2RegWeight=l2), np.zeros(d), ds)
    w_cpu, f_cpu = cpu_opt.optimize(OracleGLMObjective(loss, l2RegWeight=l2), np.zeros(d), ref)
    assert gpu_opt.convergenceReason == cpu_opt.convergenceReason != "MaxIterations"
    assert gpu_opt.currentState.iter == cpu_opt.currentState.iter
    assert gpu_opt.cg_iterations == cpu_opt.cg_iterations
    assert gpu_opt.hessian_vector_evaluations == cpu_opt.hessian_vector_evaluations > 0
    assert abs(f_gpu - f_cpu) <= 1e-9 * abs(f_cpu)
    assert np.max(np.abs(w_gpu - w_cpu)) <= 1e-6 * max(1.0, np.max(np.abs(w_cpu)))
    ds.close()


def test_owlqn_l1_logistic_with_normalization(ctx):
    # configs[2] in miniature: OWLQN with a constant L1 weight, STANDARDIZATION context on heart.txt
    import photon_ml_b200 as pml
    from photon_ml_b200.function import DistributedGLMLossFunction, NormalizationContext
    from photon_ml_b200.optimization import OWLQN
    n, d, row_ptr, col, val, label = util.read_libsvm(os.path.join(util.GOLDEN, "heart.txt"), d=13, add_intercept=True)
    f, s, icpt = util.normalization_context("STANDARDIZATION", n, d, row_ptr, col, val, d - 1)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
    ref = oracle.Data(n, d, row_ptr, col, val, label)
    ctx_n = NormalizationContext(f, s, icpt)
    gpu_opt, cpu_opt = OWLQN(10.0, ctx_n), OWLQN(10.0, ctx_n)
    w_gpu, f_gpu = gpu_opt.optimize(DistributedGLMLossFunction.create(pml.LOGISTIC), np.zeros(d), ds)
    w_cpu, f_cpu = cpu_opt.optimize(OracleGLMObjective(oracle.LOGISTIC), np.zeros(d), ref)
    assert gpu_opt.currentState.iter == cpu_opt.currentState.iter
    assert np.count_nonzero(w_gpu) == np.count_nonzero(w_cpu) < d        # L1 zeroed some coefficients
    assert np.max(np.abs(w_gpu - w_cpu)) <= 1e-6 * max(1.0, np.max(np.abs(w_cpu)))
    ds.close()


def test_tron_with_device_resident_state_matches_host_state(ctx):
    # the same TRON driver with torch tensors on the GPU (the *_dev entry points) and with numpy vectors
    import torch
    import photon_ml_b200 as pml
    from photon_ml_b200.function import DeviceGLMLossFunction, DistributedGLMLossFunction, NoNormalization
    from photon_ml_b200.optimization import TRON
    n, d, row_ptr, col, val, label = util.read_libsvm(os.path.join(util.GOLDEN, "a9a_1605.libsvm"), d=123, add_intercept=True)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
    host_opt, dev_opt = TRON(NoNormalization()), TRON(NoNormalization())
    w_host, f_host = host_opt.optimize(DistributedGLMLossFunction.create(pml.LOGISTIC, l2RegWeight=10.0), np.zeros(d), ds)
    obj = DeviceGLMLossFunction(ctx, pml.LOGISTIC, l2RegWeight=10.0)
    with obj.stream():
        w_dev, f_dev = dev_opt.optimize(obj, torch.zeros(d, dtype=torch.float64, device="cuda:0"), ds)
        torch.cuda.synchronize()
    assert dev_opt.currentState.iter == host_opt.currentState.iter
    assert dev_opt.cg_iterations == host_opt.cg_iterations
    assert abs(f_dev - f_host) <= 1e-9 * abs(f_host)
    assert np.max(np.abs(w_dev.cpu().numpy() - w_host)) <= 1e-6 * max(1.0, np.max(np.abs(w_host)))
    ds.close()


def test_compute_variances_matches_the_closed_form(ctx):
    """computeVariances = 1 / (diag(H) + EPSILON) with the L2 term in the diagonal; for the squared loss the diagonal
    is sum_i wt_i x_ij^2 + lambda whatever the coefficients (DistributedOptimizationProblemTest.scala:280-373)."""
    import photon_ml_b200 as pml
    from photon_ml_b200.function import DistributedGLMLossFunction, DistributedSmoothedHingeLossFunction
    from photon_ml_b200.optimization import compute_variances
    rng = np.random.default_rng(17)
    n, d, lam = 800, 60, 0.7
    row_ptr, col, val = util.random_csr(rng, n, d, 6)
    weight = rng.uniform(0.5, 2.0, size=n)
    label = rng.normal(size=n)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label, None, weight)
    w = rng.normal(size=d)
    got = compute_variances(DistributedGLMLossFunction.create(pml.SQUARED, l2RegWeight=lam), ds, w)
    X = util.csr_to_dense(n, d, row_ptr, col, val)
    expect = 1.0 / ((weight[:, None] * X * X).sum(axis=0) + lam + 1e-12)
    assert util.rel_err(got, expect) <= 1e-11
    ref = oracle.Data(n, d, row_ptr, col, val, (label > 0).astype(np.float64), None, weight)
    ds.set_labels(ref.label)
    got = compute_variances(DistributedGLMLossFunction.create(pml.LOGISTIC, l2RegWeight=lam), ds, w)
    assert util.rel_err(got, 1.0 / (oracle.hessian_diagonal(ref, oracle.LOGISTIC, w) + lam + 1e-12)) <= 1e-11
    assert compute_variances(DistributedSmoothedHingeLossFunction.create(), ds, w) is None
    ds.close()
