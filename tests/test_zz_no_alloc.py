"""The allocation-free objective value (pml_value_noalloc_dev) against the packed-data-set path.

Collected last on purpose (file name): if it fails under `-x`, every other test has already run.
The entry point serves callers that cannot let cudaMalloc run -- a JNI shim inside a
GetPrimitiveArrayCritical region, executor threads sharing one GPU -- so it works from the
caller's own device arrays and a fixed scratch array inside the library.  Here it is called twice
in a row (second time with other coefficients and offsets, so anything stale in the shared scratch
would show) and compared with pml_value_and_gradient on a data set packed from the same arrays
and with the CPU oracle, to the tolerance of the other GPU parity tests; the amount of free device
memory must be the same after the calls as before them (that is what the entry point is for).
"""
import numpy as np
import pytest

import oracle
from tests import util

pytestmark = pytest.mark.gpu

TOL = 1e-11  # same as tests/test_gpu_parity.py


def test_value_noalloc_matches_packed_path_on_two_consecutive_calls():
    import torch
    import photon_ml_b200 as pml

    rng = np.random.default_rng(123)
    n, d = 70_000, 5_000   # 274 CTAs: the cross-CTA reduction and the completion counter are exercised
    row_ptr, col, val = util.fast_random_csr(rng, n, d, 12, skew=1.1)
    label = util.labels_for(rng, 0, n)
    weight = rng.uniform(0.5, 2.0, size=n)
    dev = torch.device("cuda", 0)
    t_rp = torch.tensor(row_ptr, device=dev)
    t_col = torch.tensor(col, device=dev)
    t_val = torch.tensor(val.astype(np.float32), device=dev)
    t_lab = torch.tensor(label, device=dev)
    t_wt = torch.tensor(weight, device=dev)
    out = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)

    ctx = pml.Context(0)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label, None, weight)
    val32 = val.astype(np.float32).astype(np.float64)
    # the start-up call (resolves the scratch address, loads the kernels): the one call that may allocate
    t_zero = torch.zeros(d, dtype=torch.float64, device=dev)
    pml.abi.value_noalloc_dev(0, None, n, d, t_rp.data_ptr(), t_col.data_ptr(), t_val.data_ptr(), pml.abi.F32,
                              t_lab.data_ptr(), None, t_wt.data_ptr(), pml.LOGISTIC, t_zero.data_ptr(), out.data_ptr())
    torch.cuda.synchronize()
    for call, loss in enumerate((pml.LOGISTIC, pml.POISSON)):
        coef = rng.normal(scale=0.2, size=d)
        offset = rng.normal(scale=0.3, size=n) if call == 1 else None
        ds.set_offsets(offset)
        want, _ = ds.value_and_gradient(loss, coef, want_gradient=False)
        want_cpu, _ = oracle.value_and_gradient(oracle.Data(n, d, row_ptr, col, val32, label, offset, weight), loss, coef)
        t_coef = torch.tensor(coef, device=dev)
        t_off = None if offset is None else torch.tensor(offset, device=dev)
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info(dev)
        current_before = torch.cuda.current_device()
        pml.abi.value_noalloc_dev(0, None, n, d, t_rp.data_ptr(), t_col.data_ptr(), t_val.data_ptr(), pml.abi.F32,
                                  t_lab.data_ptr(), None if t_off is None else t_off.data_ptr(), t_wt.data_ptr(), loss,
                                  t_coef.data_ptr(), out.data_ptr())
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info(dev)
        assert free_after == free_before, (call, free_before, free_after)   # nothing was allocated or freed
        assert torch.cuda.current_device() == current_before
        got = out.item()
        assert abs(got - want) <= TOL * abs(want), (call, got, want)
        assert abs(got - want_cpu) <= TOL * abs(want_cpu), (call, got, want_cpu)

    with pytest.raises(pml.IllegalArgumentError):  # beyond the fixed scratch: refused, nothing launched
        pml.abi.value_noalloc_dev(0, None, 16_777_217, d, t_rp.data_ptr(), t_col.data_ptr(), t_val.data_ptr(),
                                  pml.abi.F32, t_lab.data_ptr(), None, None, pml.LOGISTIC, t_coef.data_ptr(),
                                  out.data_ptr())
    ds.close()
    ctx.close()
