"""One-off measurement (not part of bench.py): BASELINE.json configs[4], the GAME ads-CTR shape, on ONE B200 --
fixed effect over 40 M rows x 25 indicator features (1e9 stored entries, d = 2 M, Zipf column popularity) + a
per-member random effect with 10 M entities (power-law rows per entity, mean 4; 10 indicator features per row in a
32-dimensional projected space), logistic loss, L2 = 1, coordinate descent fixed -> per-member, everything per-row in
device memory (photon_ml_b200/game.py).  SCALE=0.1 shrinks rows and entities."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import photon_ml_b200 as pml  # noqa: E402
from photon_ml_b200.game import CoordinateDescent, FixedEffect, RandomEffect  # noqa: E402

SCALE = float(os.environ.get("SCALE", "1.0"))
N_MEMBERS, D_FE, K_FE, K_RE, D_RE = int(10_000_000 * SCALE), 2_000_000, 25, 10, 32
PASSES = int(os.environ.get("PASSES", "2"))


def main():
    dev = torch.device("cuda", 0)
    ctx = pml.Context(0)
    gen = torch.Generator(device=dev)
    gen.manual_seed(5)
    t0 = time.time()
    # power-law rows per member with mean ~4: Pareto tail, capped
    u = torch.rand(N_MEMBERS, generator=gen, device=dev, dtype=torch.float64)
    rows = torch.clamp((1.0 / u.pow(1.0 / 1.35)).floor(), max=5000).to(torch.int64)
    n = int(rows.sum())
    ent_row_ptr = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), torch.cumsum(rows, 0)])
    p = torch.arange(1, D_FE + 1, device=dev, dtype=torch.float64).pow(-1.05)
    cdf = torch.cumsum(p / p.sum(), 0)
    fe_cols = torch.empty((n, K_FE), dtype=torch.int32, device=dev)
    step = 4_000_000
    for a in range(0, n, step):
        b = min(n, a + step)
        c = torch.searchsorted(cdf, torch.rand((b - a, K_FE), generator=gen, device=dev, dtype=torch.float64)).clamp_(max=D_FE - 1)
        fe_cols[a:b] = torch.sort(c, dim=1)[0].to(torch.int32)
    re_cols = torch.sort(torch.randint(0, D_RE, (n, K_RE), generator=gen, device=dev, dtype=torch.int32), dim=1)[0]
    label = (torch.rand(n, generator=gen, device=dev) < 0.03).to(torch.float64)
    torch.cuda.synchronize()
    print("generated %d rows for %d members (mean %.2f rows, largest %d) in %.1f s" %
          (n, N_MEMBERS, n / N_MEMBERS, int(rows.max()), time.time() - t0), flush=True)

    t0 = time.time()
    fe_ptr = torch.arange(n + 1, dtype=torch.int64, device=dev) * K_FE
    ones = torch.ones(n * K_FE, dtype=torch.float32, device=dev)
    fe_ds = pml.Dataset.from_device_pointers(ctx, n, D_FE, fe_ptr.data_ptr(), fe_cols.data_ptr(), ones.data_ptr(), pml.abi.F32,
                                             label.data_ptr())
    ctx.synchronize()
    t_fe = time.time() - t0
    del fe_cols, fe_ptr, ones
    torch.cuda.empty_cache()
    t0 = time.time()
    batch = pml.Batch(ctx, ent_row_ptr.cpu().numpy(), np.full(N_MEMBERS, D_RE, np.int32), np.arange(n + 1, dtype=np.int64) * K_RE,
                      re_cols.cpu().numpy().ravel(), np.ones(n * K_RE, np.float32), label.cpu().numpy())
    ctx.synchronize()
    del re_cols
    torch.cuda.empty_cache()
    print("packed: fixed effect %.1f s, %.2f GB (%d B/value); per-member batch %.1f s (host arrays), %.2f GB" %
          (t_fe, fe_ds.info().device_bytes / 1e9, fe_ds.info().value_bytes, time.time() - t0,
           batch.dataset.info().device_bytes / 1e9), flush=True)

    coords = {"fixed": FixedEffect(fe_ds, pml.LOGISTIC, l2RegWeight=1.0),
              "per-member": RandomEffect(batch, None, pml.LOGISTIC, l2RegWeight=1.0)}
    cd = CoordinateDescent(coords, n, descent_iterations=PASSES)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    _, history = cd.run()
    wall = time.perf_counter() - t0
    for h in history:
        print(json.dumps(h))
    print("%d coordinate-descent passes over %d rows, %d members: %.2f s; peak device memory %.1f GB" %
          (PASSES, n, N_MEMBERS, wall, (torch.cuda.mem_get_info()[1] - torch.cuda.mem_get_info()[0]) / 1e9))


if __name__ == "__main__":
    main()
