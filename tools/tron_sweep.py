"""TRON on the bench workload (configs[1]) with device-resident state for several L2 weights: which settings reach
a convergence reason other than the iteration cap, and how long that takes."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
import photon_ml_b200 as pml  # noqa: E402
from photon_ml_b200.function import DeviceGLMLossFunction, NoNormalization  # noqa: E402
from photon_ml_b200.optimization import TRON  # noqa: E402

dev = torch.device("cuda", 0)
ctx = pml.Context(0)
rows = bench.generate_rows(torch, dev, range(bench.N_ROWS // bench.GEN_CHUNK))
ds = bench.pack_rows(pml, ctx, torch, rows)
del rows
for lam in (1.0, 10.0, 100.0, 1000.0):
    obj = DeviceGLMLossFunction(ctx, bench.LOSS_POISSON, l2RegWeight=lam)
    tron = TRON(NoNormalization(), maxNumIterations=100)
    with obj.stream():
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        _, f = tron.optimize(obj, torch.zeros(bench.N_FEATS, dtype=torch.float64, device=dev), ds)
        torch.cuda.synchronize()
    print("lambda %g: %.2f s, %d iterations, %s, %d value+gradient, %d Hessian-vector, objective %.6f" % (
        lam, time.perf_counter() - t0, tron.currentState.iter, tron.convergenceReason, tron.evaluations,
        tron.hessian_vector_evaluations, f), flush=True)
