"""The lock-step batched L-BFGS / OWL-QN (photon_ml_b200/batched_lbfgs.py) against one run of the single-problem
driver (optimization.LBFGS / OWLQN) per entity on the CPU oracle objective -- the optimizer GAME's coordinates
default to (api/optimization/OptimizerFactory.scala:46-62), run entity by entity by RandomEffectCoordinate.updateModel
(api/algorithm/RandomEffectCoordinate.scala:108-119).  `-m gpu`."""
import numpy as np
import pytest

import oracle
from tests import util
from tests.objectives import OracleGLMObjective
from tests.test_gpu_batch import _entities

pytestmark = pytest.mark.gpu


def _check(loss, l2, l1, n_ent=120):
    import photon_ml_b200 as pml
    from photon_ml_b200.batched import reason_names
    from photon_ml_b200.batched_lbfgs import BatchedLBFGS
    from photon_ml_b200.function import NoNormalization
    from photon_ml_b200.optimization import LBFGS, OWLQN
    rng = np.random.default_rng(70 + loss + int(10 * l1))
    ent_row_ptr, dims, row_ptr, col, val, n = _entities(rng, n_ent, 12, 25)
    label = util.labels_for(rng, loss, n)
    offset = rng.normal(scale=0.2, size=n)
    weight = rng.uniform(0.5, 2.0, size=n)
    ctx = pml.Context(0)
    batch = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, col, val, label, offset, weight)
    solver = BatchedLBFGS(batch, loss, l2RegWeight=l2, l1RegWeight=l1)
    w, f, info = solver.optimize()
    w, f = w.cpu().numpy(), f.cpu().numpy()
    reasons = reason_names(info["reasons"])
    mismatched = 0
    total_single = 0
    for e in range(len(dims)):
        lo, hi = int(ent_row_ptr[e]), int(ent_row_ptr[e + 1])
        a, b = int(batch.coef_ptr[e]), int(batch.coef_ptr[e + 1])
        sl = slice(int(row_ptr[lo]), int(row_ptr[hi]))
        local = oracle.Data(hi - lo, int(dims[e]), row_ptr[lo:hi + 1] - row_ptr[lo], col[sl], val[sl], label[lo:hi],
                            offset[lo:hi], weight[lo:hi])
        opt = OWLQN(l1, NoNormalization()) if l1 > 0 else LBFGS(NoNormalization())
        w_ref, f_ref = opt.optimize(OracleGLMObjective(loss, l2RegWeight=l2), np.zeros(int(dims[e])), local)
        total_single += opt.evaluations
        dw = np.max(np.abs(w[a:b] - w_ref), initial=0.0) / max(1.0, np.max(np.abs(w_ref), initial=0.0))
        df = abs(f[e] - f_ref) / max(abs(f_ref), 1e-300)
        same_path = (info["iterations"][e], reasons[e], info["entity_evaluations"][e]) == \
            (opt.currentState.iter, opt.convergenceReason, opt.evaluations)
        if not (same_path and dw <= 1e-6 and df <= 1e-9):
            # an entity whose line search, orthant projection or stopping test sat exactly on a threshold took another
            # (equally valid) path: it still ends at the same optimum, to the optimizer's own tolerance
            mismatched += 1
            assert dw <= 1e-3 and df <= 1e-4, (e, dw, df)
    # same operations in the same order; the summation order inside the dot products differs in the last bits, which
    # can flip a line-search or convergence test that sits exactly on its threshold for an odd entity
    assert mismatched <= max(2, n_ent // 40), mismatched
    # one batched call serves every entity that is still searching: far fewer calls than per-entity evaluations
    assert info["evaluations"] < total_single / 8
    batch.close()
    ctx.close()


@pytest.mark.parametrize("loss,l2", [(oracle.LOGISTIC, 1.0), (oracle.POISSON, 5.0), (oracle.SQUARED, 0.5),
                                     (oracle.SMOOTHED_HINGE, 1.0)])
def test_every_entity_follows_its_own_lbfgs_run(loss, l2):
    _check(loss, l2, 0.0)


@pytest.mark.parametrize("loss,l1", [(oracle.LOGISTIC, 0.5), (oracle.SQUARED, 0.3)])
def test_every_entity_follows_its_own_owlqn_run(loss, l1):
    # (a little L2 next to the L1 term keeps every tiny per-entity problem well conditioned: without it an entity whose
    # rows barely determine a coefficient ends 1e-4 apart in that coefficient at an objective difference of 1e-9)
    _check(loss, 0.1, l1)


@pytest.mark.parametrize("max_dim,n_ent", [(20, 300), (150, 200), (1500, 40)])
@pytest.mark.parametrize("l1,m", [(0.0, 10), (0.3, 10), (0.0, 3)])
def test_fused_direction_is_the_unfused_recursion_rounding_for_rounding(max_dim, n_ent, l1, m):
    """pml_batch_lbfgs_direction_dev / pml_batch_lbfgs_push_dev (one kernel for the two-loop recursion over per-entity
    rings of correction pairs) against the same algebra as 2m dots and 2m + 1 axpys over shifted lists: bit-identical
    iterates.  max_dim 150: 32 consecutive entities do not fit the staging area in one go, and entities on both sides
    of the 64-coefficient switch to whole-warp dots; 1500: entities larger than the staging area.  m = 3: rings wrap
    early and the oldest pairs are overwritten."""
    import photon_ml_b200 as pml
    from photon_ml_b200.batched_lbfgs import BatchedLBFGS
    rng = np.random.default_rng(17)
    ent_row_ptr, dims, row_ptr, col, val, n = _entities(rng, n_ent, max_dim, 30)
    label = util.labels_for(rng, oracle.LOGISTIC, n)
    ctx = pml.Context(0)
    batch = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, col, val, label, rng.normal(scale=0.2, size=n))
    out = []
    for fused in (True, False):
        solver = BatchedLBFGS(batch, oracle.LOGISTIC, l2RegWeight=0.1, l1RegWeight=l1, numCorrections=m,
                              maxNumIterations=40, fused_direction=fused)
        assert solver.fused_direction == fused
        w, f, info = solver.optimize()
        out.append((w.cpu().numpy(), f.cpu().numpy(), info["iterations"], info["entity_evaluations"], info["reasons"],
                    info["evaluations"]))
    a, b = out
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3]) and np.array_equal(a[4], b[4]) and a[5] == b[5]
    assert a[2].max() > m            # the rings did wrap
    batch.close()
    ctx.close()


def test_more_than_16_correction_pairs_take_the_unfused_recursion():
    import photon_ml_b200 as pml
    from photon_ml_b200.batched_lbfgs import BatchedLBFGS
    rng = np.random.default_rng(18)
    ent_row_ptr, dims, row_ptr, col, val, n = _entities(rng, 30, 12, 20)
    ctx = pml.Context(0)
    batch = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, col, val, util.labels_for(rng, oracle.LOGISTIC, n))
    solver = BatchedLBFGS(batch, oracle.LOGISTIC, l2RegWeight=0.1, numCorrections=20)
    assert not solver.fused_direction
    w, f, info = solver.optimize()
    assert np.all(np.isfinite(w.cpu().numpy()))
    with pytest.raises(pml.IllegalArgumentError):
        batch.lbfgs_direction_dev(None, 17, w.data_ptr(), w.data_ptr(), w.data_ptr(), w.data_ptr(), w.data_ptr(),
                                  w.data_ptr(), w.data_ptr())
    batch.close()
    ctx.close()
