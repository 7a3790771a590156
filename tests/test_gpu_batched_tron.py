"""The lock-step batched TRON (photon_ml_b200/batched.py) against one run of the single-problem TRON restatement
per entity on the CPU oracle objective: what RandomEffectCoordinate.updateModel does entity by entity
(api/algorithm/RandomEffectCoordinate.scala:108-119).  `-m gpu`."""
import numpy as np
import pytest

import oracle
from tests import util
from tests.objectives import OracleGLMObjective
from tests.test_gpu_batch import _entities

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("loss,l2", [(oracle.LOGISTIC, 1.0), (oracle.POISSON, 5.0), (oracle.SQUARED, 0.5)])
def test_every_entity_follows_its_own_tron_run(loss, l2):
    import photon_ml_b200 as pml
    from photon_ml_b200.batched import BatchedTRON, reason_names
    from photon_ml_b200.function import NoNormalization
    from photon_ml_b200.optimization import TRON
    rng = np.random.default_rng(40 + loss)
    ent_row_ptr, dims, row_ptr, col, val, n = _entities(rng, 160, 12, 25)
    label = util.labels_for(rng, loss, n)
    offset = rng.normal(scale=0.2, size=n)
    weight = rng.uniform(0.5, 2.0, size=n)
    ctx = pml.Context(0)
    batch = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, col, val, label, offset, weight)
    solver = BatchedTRON(batch, loss, l2RegWeight=l2)
    w, f, info = solver.optimize()
    w, f = w.cpu().numpy(), f.cpu().numpy()
    reasons = reason_names(info["reasons"])
    assert info["evaluations"] < 60 and info["hessian_vector_evaluations"] < 15 * 20 * 5   # batched calls, not per entity
    mismatched_counts = 0
    for e in range(len(dims)):
        lo, hi = int(ent_row_ptr[e]), int(ent_row_ptr[e + 1])
        a, b = int(batch.coef_ptr[e]), int(batch.coef_ptr[e + 1])
        sl = slice(int(row_ptr[lo]), int(row_ptr[hi]))
        local = oracle.Data(hi - lo, int(dims[e]), row_ptr[lo:hi + 1] - row_ptr[lo], col[sl], val[sl], label[lo:hi],
                            offset[lo:hi], weight[lo:hi])
        opt = TRON(NoNormalization())
        w_ref, f_ref = opt.optimize(OracleGLMObjective(loss, l2RegWeight=l2), np.zeros(int(dims[e])), local)
        assert np.max(np.abs(w[a:b] - w_ref), initial=0.0) <= 1e-6 * max(1.0, np.max(np.abs(w_ref), initial=0.0)), e
        assert abs(f[e] - f_ref) <= 1e-9 * max(abs(f_ref), 1e-300), e
        if (info["iterations"][e], reasons[e], info["cg_iterations"][e]) != \
                (opt.currentState.iter, opt.convergenceReason, sum(opt.cg_iterations)):
            mismatched_counts += 1
    # same operations in the same order; only the summation order inside the dot products differs in the last
    # bits, which can flip a convergence test that sits exactly on its threshold for an odd entity
    assert mismatched_counts <= 2
    batch.close()
    ctx.close()


def test_segment_dot_and_axpby():
    import torch
    import photon_ml_b200 as pml
    rng = np.random.default_rng(3)
    ent_row_ptr, dims, row_ptr, col, val, n = _entities(rng, 300, 90, 3)
    ctx = pml.Context(0)
    batch = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, col, val, np.zeros(n))
    E, D = batch.n_entities, batch.d_total
    dev = torch.device("cuda", 0)
    x, y = rng.normal(size=D), rng.normal(size=D)
    a, c = rng.normal(size=E), rng.normal(size=E)
    active = rng.random(E) < 0.7
    with torch.cuda.stream(torch.cuda.ExternalStream(ctx.stream, device=dev)):
        tx, ty = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
        ta, tc = torch.from_numpy(a).to(dev), torch.from_numpy(c).to(dev)
        tact = torch.from_numpy(active.astype(np.uint8)).to(dev)
        out_e = torch.empty(E, dtype=torch.float64, device=dev)
        out_d = torch.full((D,), 7.0, dtype=torch.float64, device=dev)
        batch.segment_dot_dev(tx.data_ptr(), ty.data_ptr(), out_e.data_ptr())
        batch.segment_axpby_dev(ta.data_ptr(), tx.data_ptr(), tc.data_ptr(), ty.data_ptr(), out_d.data_ptr(), tact.data_ptr())
        ctx.synchronize()
        dots, res = out_e.cpu().numpy(), out_d.cpu().numpy()
        out_d2 = tx.clone()
        batch.segment_axpby_dev(ta.data_ptr(), out_d2.data_ptr(), None, ty.data_ptr(), out_d2.data_ptr())   # in place, c = 1
        ctx.synchronize()
        res2 = out_d2.cpu().numpy()
    ent = np.repeat(np.arange(E), dims)
    for e in range(E):
        s = slice(int(batch.coef_ptr[e]), int(batch.coef_ptr[e + 1]))
        assert abs(dots[e] - float(np.dot(x[s], y[s]))) <= 1e-12 * max(1.0, float(np.abs(x[s] * y[s]).sum()))
    expect = np.where(active[ent], a[ent] * x + c[ent] * y, 7.0)
    assert np.array_equal(res, expect)          # one rounding per product, one per sum: the same as numpy's
    assert np.array_equal(res2, a[ent] * x + y)
    batch.close()
    ctx.close()


@pytest.mark.parametrize("max_dim,n_ent", [(20, 400), (150, 300), (1500, 70)])
def test_fused_cg_step_is_the_unfused_loop_rounding_for_rounding(max_dim, n_ent):
    """pml_batch_cg_step_dev does an iteration's per-entity vector algebra in one kernel, with the unfused loop's
    rounding: a lane per entity below 64 coefficients (dots in index order), the whole warp with the fixed shuffle
    tree from 64 on.  max_dim 150: 32 consecutive entities do not fit the staging area in one go; 1500: single
    entities larger than the staging area."""
    import photon_ml_b200 as pml
    from photon_ml_b200.batched import BatchedTRON
    rng = np.random.default_rng(91)
    ent_row_ptr, dims, row_ptr, col, val, n = _entities(rng, n_ent, max_dim, 30)
    label = util.labels_for(rng, oracle.LOGISTIC, n)
    ctx = pml.Context(0)
    batch = pml.Batch(ctx, ent_row_ptr, dims, row_ptr, col, val, label, rng.normal(scale=0.2, size=n))
    out = {}
    for fused in (True, False):
        for l2 in (0.0, 0.7):
            w, f, info = BatchedTRON(batch, oracle.LOGISTIC, l2RegWeight=l2, fused_cg=fused).optimize()
            out[(fused, l2)] = (w.cpu().numpy(), f.cpu().numpy(), info["iterations"], info["cg_iterations"], info["reasons"])
    for l2 in (0.0, 0.7):
        a, b = out[(True, l2)], out[(False, l2)]
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        assert np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3]) and np.array_equal(a[4], b[4])
    batch.close()
    ctx.close()
