"""One-off measurement (not part of bench.py): BASELINE.json configs[3], the GLMix MovieLens-20M-shaped problem --
fixed effect (d = 1000, 30 indicator features per row) + per-user random effect (138 493 users, 5 of 20 movie-side
indicators + intercept) + per-movie random effect (27 278 movies, 6 of 30 user-side indicators + intercept), logistic
loss, L2 = 1, 3 coordinate-descent passes in the order fixed -> per-user -> per-movie, everything per-row kept in
device memory (photon_ml_b200/game.py).  SCALE=0.1 shrinks the row count for a quick look."""
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
N_USERS, N_MOVIES, D_FE = int(138_493 * SCALE), int(27_278 * SCALE), 1000


def indicators(gen, dev, n, k, d, intercept):
    cols = torch.randint(0, d, (n, k), generator=gen, device=dev, dtype=torch.int32)
    cols, _ = torch.sort(cols, dim=1)
    if intercept:
        cols = torch.cat([cols, torch.full((n, 1), d, device=dev, dtype=torch.int32)], dim=1)
    return cols


def main():
    dev = torch.device("cuda", 0)
    ctx = pml.Context(0)
    gen = torch.Generator(device=dev)
    gen.manual_seed(4)
    t0 = time.time()
    rows = torch.exp(torch.randn(N_USERS, generator=gen, device=dev, dtype=torch.float64) + np.log(144.0) - 0.5)
    rows = rows.clamp_(min=20).to(torch.int64)
    n = int(rows.sum())
    user_row_ptr = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), torch.cumsum(rows, 0)])
    p = torch.arange(1, N_MOVIES + 1, device=dev, dtype=torch.float64).pow(-0.9)
    cdf = torch.cumsum(p / p.sum(), 0)
    movie_of_row = torch.searchsorted(cdf, torch.rand(n, generator=gen, device=dev, dtype=torch.float64)).clamp_(max=N_MOVIES - 1)
    movie_map = torch.argsort(movie_of_row, stable=True)                 # per-movie row i = global row movie_map[i]
    movie_rows = torch.bincount(movie_of_row, minlength=N_MOVIES)
    movie_row_ptr = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), torch.cumsum(movie_rows, 0)])
    label = (torch.rand(n, generator=gen, device=dev) < 0.5).to(torch.float64)
    fe_cols = indicators(gen, dev, n, 30, D_FE, False)
    user_cols = indicators(gen, dev, n, 5, 20, True)                     # rows already grouped by user
    movie_cols = indicators(gen, dev, n, 6, 30, True)                    # generated in per-movie order
    torch.cuda.synchronize()
    print("generated %d rows, %d users, %d movies (largest movie %d rows) in %.1f s" %
          (n, N_USERS, N_MOVIES, int(movie_rows.max()), time.time() - t0), flush=True)

    t0 = time.time()
    fe_ptr = torch.arange(n + 1, dtype=torch.int64, device=dev) * 30
    ones = torch.ones(n * 30, dtype=torch.float32, device=dev)
    fe_ds = pml.Dataset.from_device_pointers(ctx, n, D_FE, fe_ptr.data_ptr(), fe_cols.data_ptr(), ones.data_ptr(), pml.abi.F32,
                                             label.data_ptr())
    ctx.synchronize()
    del fe_cols, fe_ptr
    label_h = label.cpu().numpy()
    user_batch = pml.Batch(ctx, user_row_ptr.cpu().numpy(), np.full(N_USERS, 21, np.int32), np.arange(n + 1, dtype=np.int64) * 6,
                           user_cols.cpu().numpy().ravel(), np.ones(n * 6, np.float32), label_h)
    del user_cols
    movie_batch = pml.Batch(ctx, movie_row_ptr.cpu().numpy(), np.full(N_MOVIES, 31, np.int32),
                            np.arange(n + 1, dtype=np.int64) * 7, movie_cols.cpu().numpy().ravel(), np.ones(n * 7, np.float32),
                            label_h[movie_map.cpu().numpy()])
    del movie_cols, ones
    torch.cuda.empty_cache()
    ctx.synchronize()
    print("packed in %.1f s: fixed effect %.2f GB (%d B/value), per-user %.2f GB, per-movie %.2f GB" %
          (time.time() - t0, fe_ds.info().device_bytes / 1e9, fe_ds.info().value_bytes,
           user_batch.dataset.info().device_bytes / 1e9, movie_batch.dataset.info().device_bytes / 1e9), flush=True)

    coords = {"fixed": FixedEffect(fe_ds, pml.LOGISTIC, l2RegWeight=1.0),
              "per-user": RandomEffect(user_batch, None, pml.LOGISTIC, l2RegWeight=1.0),
              "per-movie": RandomEffect(movie_batch, movie_map, pml.LOGISTIC, l2RegWeight=1.0)}
    cd = CoordinateDescent(coords, n, descent_iterations=3)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    _, history = cd.run()
    wall = time.perf_counter() - t0
    for h in history:
        print(json.dumps(h))
    print("3 coordinate-descent passes over %d rows: %.2f s" % (n, wall))


if __name__ == "__main__":
    main()
