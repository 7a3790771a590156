"""State that the objective classes keep between calls -- which NormalizationContext is set on a packed data set,
which coefficient vector the cached curvature of a Hessian-vector product belongs to -- when several objectives take
turns on ONE data set.  The reference passes the context and the coefficients on every call
(DistributedGLMLossFunction.scala:93-137), so nothing may depend on who called last.  `-m gpu`."""
import numpy as np
import pytest

import oracle
from tests import util

pytestmark = pytest.mark.gpu

TOL = 1e-11


def _problem(seed=5):
    rng = np.random.default_rng(seed)
    n, d = 900, 41
    row_ptr, col, val = util.random_csr(rng, n, d - 1, 6, skew=1.1)
    row_ptr, col, val = util.append_intercept(row_ptr, col, val, d - 1)
    label = util.labels_for(rng, oracle.LOGISTIC, n)
    return rng, n, d, row_ptr, col, val, label


def test_objectives_taking_turns_on_one_dataset_each_get_their_own_context():
    import photon_ml_b200 as pml
    from photon_ml_b200.function import DistributedGLMLossFunction, NoNormalization, NormalizationContext
    rng, n, d, row_ptr, col, val, label = _problem()
    ctx = pml.Context(0)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
    ref = oracle.Data(n, d, row_ptr, col, val, label)
    fa, sa, icpt = util.normalization_context("STANDARDIZATION", n, d, row_ptr, col, val, d - 1)
    fb, _, _ = util.normalization_context("SCALE_WITH_MAX_MAGNITUDE", n, d, row_ptr, col, val, d - 1)
    ctx_a, ctx_b, none = NormalizationContext(fa, sa, icpt), NormalizationContext(fb, None, icpt), NoNormalization()
    obj_a, obj_b, obj_c = (DistributedGLMLossFunction(pml.LOGISTIC) for _ in range(3))
    w, v = rng.normal(scale=0.3, size=d), rng.normal(size=d)
    want = {"a": oracle.value_and_gradient(ref, oracle.LOGISTIC, w, fa, sa, icpt),
            "b": oracle.value_and_gradient(ref, oracle.LOGISTIC, w, fb, None, icpt),
            "c": oracle.value_and_gradient(ref, oracle.LOGISTIC, w)}
    for who in "abacbcab":        # A skips nothing because IT applied ctx_a last time: the data set remembers, not A
        obj, nc = {"a": (obj_a, ctx_a), "b": (obj_b, ctx_b), "c": (obj_c, none)}[who]
        f, g = obj.calculate(ds, w, nc)
        assert abs(f - want[who][0]) <= TOL * abs(want[who][0]), who
        assert util.rel_err(g, want[who][1]) <= TOL, who
    hv = obj_c.hessianVector(ds, w, v, none)      # an empty context clears what B left behind
    assert util.rel_err(hv, oracle.hessian_vector(ref, oracle.LOGISTIC, w, v)) <= TOL
    # a NEW data set never inherits another one's record, whatever its id() is
    ds.close()
    ds2 = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
    f, g = obj_a.calculate(ds2, w, ctx_a)
    assert abs(f - want["a"][0]) <= TOL * abs(want["a"][0]) and util.rel_err(g, want["a"][1]) <= TOL
    ds2.close()
    ctx.close()


def test_device_objective_contexts_and_curvature_tags():
    import torch
    import photon_ml_b200 as pml
    from photon_ml_b200.function import DeviceGLMLossFunction, NoNormalization, NormalizationContext
    rng, n, d, row_ptr, col, val, label = _problem(6)
    ctx = pml.Context(0)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
    ref = oracle.Data(n, d, row_ptr, col, val, label)
    fa, sa, icpt = util.normalization_context("STANDARDIZATION", n, d, row_ptr, col, val, d - 1)
    nc = NormalizationContext(fa, sa, icpt)
    dev = torch.device("cuda", 0)
    a, b = DeviceGLMLossFunction(ctx, pml.LOGISTIC), DeviceGLMLossFunction(ctx, pml.LOGISTIC)
    with a.stream():
        w = torch.from_numpy(rng.normal(scale=0.3, size=d)).to(dev)
        v = torch.from_numpy(rng.normal(size=d)).to(dev)
        f, g = a.calculate(ds, w, nc)
        f_ref, g_ref = oracle.value_and_gradient(ref, oracle.LOGISTIC, w.cpu().numpy(), fa, sa, icpt)
        assert abs(f - f_ref) <= TOL * abs(f_ref) and util.rel_err(g.cpu().numpy(), g_ref) <= TOL
        applied = ds._applied_norm
        a.calculate(ds, w, nc)
        a.hessianVector(ds, w, v, nc)
        assert ds._applied_norm is applied is nc        # the same context is not uploaded again on every call
        # the tensor updated in place is a NEW coefficient vector: the cached curvature must not be reused
        hv1 = a.hessianVector(ds, w, v, nc).cpu().numpy()
        w.add_(0.25)
        hv2 = a.hessianVector(ds, w, v, nc).cpu().numpy()
        assert util.rel_err(hv1, oracle.hessian_vector(ref, oracle.LOGISTIC, w.cpu().numpy() - 0.25, v.cpu().numpy(), fa, sa, icpt)) <= TOL
        assert util.rel_err(hv2, oracle.hessian_vector(ref, oracle.LOGISTIC, w.cpu().numpy(), v.cpu().numpy(), fa, sa, icpt)) <= TOL
        # a second objective on the same data set starts its own tags: no collision with A's
        w_b = torch.from_numpy(rng.normal(scale=0.3, size=d)).to(dev)
        hv_b = b.hessianVector(ds, w_b, v, nc).cpu().numpy()
        assert util.rel_err(hv_b, oracle.hessian_vector(ref, oracle.LOGISTIC, w_b.cpu().numpy(), v.cpu().numpy(), fa, sa, icpt)) <= TOL
        # and an empty context clears the normalization for the device objective too
        f0, _ = b.calculate(ds, w_b, NoNormalization())
        f0_ref, _ = oracle.value_and_gradient(ref, oracle.LOGISTIC, w_b.cpu().numpy())
        assert abs(f0 - f0_ref) <= TOL * abs(f0_ref)
    ds.close()
    ctx.close()


def test_coordinate_rows_without_a_score_count_as_zero():
    """game.CoordinateDescent._to_global: rows a coordinate does not cover contribute 0 to the others' offsets."""
    import torch
    from photon_ml_b200 import game

    class Coord:
        row_map = torch.tensor([0, 2, 5], device="cuda")
    cd = object.__new__(game.CoordinateDescent)
    cd.torch, cd.n, cd.device = torch, 7, torch.device("cuda", 0)
    out = cd._to_global(Coord, torch.tensor([1.0, 2.0, 3.0], dtype=torch.float64, device="cuda"))
    assert out.cpu().tolist() == [1.0, 0.0, 2.0, 0.0, 0.0, 3.0, 0.0]
