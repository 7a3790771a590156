"""Where the lock-step per-entity TRON solve spends its time: device time by kernel and the host wall time around it,
for the unfused CG loop, for the one-kernel CG step (BatchedTRON(fused_cg=True)) and for lock-step L-BFGS.  Not a bench value: torch's
profiler is attached."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import bench  # noqa: E402
import photon_ml_b200 as pml  # noqa: E402
from photon_ml_b200.batched import BatchedTRON  # noqa: E402
from torch.profiler import profile, ProfilerActivity  # noqa: E402

n_ent = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
ctx = pml.Context(0)
rng = np.random.default_rng(5)
parts = bench.make_entity_problems(rng, n_ent, 10, device=torch.device("cuda", ctx.device))
batch = pml.Batch(ctx, *parts)
BatchedTRON(batch, pml.LOGISTIC, l2RegWeight=1.0, maxNumIterations=2).optimize()
out = {}
from photon_ml_b200.batched_lbfgs import BatchedLBFGS  # noqa: E402
forms = {"unfused": lambda: BatchedTRON(batch, pml.LOGISTIC, l2RegWeight=1.0, fused_cg=False),
         "fused": lambda: BatchedTRON(batch, pml.LOGISTIC, l2RegWeight=1.0, fused_cg=True),
         "lbfgs": lambda: BatchedLBFGS(batch, pml.LOGISTIC, l2RegWeight=1.0)}
only = os.environ.get("PML_FORMS")
for name, make in forms.items():
    if only and name not in only.split(","):
        continue
    make().optimize()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    info = make().optimize()[2]
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        make().optimize()
        torch.cuda.synchronize()
    rows = []
    total = 0.0
    launches = 0
    for e in prof.key_averages():
        dev_us = getattr(e, "self_device_time_total", 0) or 0
        if dev_us > 0 and e.device_type is not None and "cuda" in str(e.device_type).lower():
            rows.append((dev_us, e.count, e.key[:90]))
            total += dev_us
            launches += e.count
    rows.sort(reverse=True)
    out[name] = {
        "wall_s_unprofiled": wall, "batched_evaluations": info["evaluations"],
        "mean_iterations_per_entity": float(info["iterations"].mean()), "device_ms_total": total / 1e3, "kernel_launches": launches,
        "top": [{"ms": r[0] / 1e3, "n": r[1], "kernel": r[2]} for r in rows[:14]]}
print(json.dumps(out, indent=1))
