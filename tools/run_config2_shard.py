"""One-off measurement (not part of the product or of bench.py): the per-GPU shard of BASELINE.json configs[2]
-- Criteo-shaped sparse logistic regression, 200 M rows x 10 M hashed features on 8 GPUs = 25 M rows per GPU,
39 entries per row (13 from a small dense range + 26 hashed Zipf(1.05) categorical), values 1.0, CTR ~3 % --
packed and evaluated on one B200.  Prints pack time, memory, evaluation times and a few self-checks."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import photon_ml_b200 as pml  # noqa: E402

N = int(os.environ.get("ROWS", 25_000_000))
D = 10_000_000
CHUNK = 500_000


def main():
    dev = torch.device("cuda", 0)
    ctx = pml.Context(0)
    g = torch.Generator(device=dev)
    g.manual_seed(3)
    ranks = torch.arange(1, D - 64 + 1, device=dev, dtype=torch.float64)
    p = ranks.pow(-1.05)
    cdf = torch.cumsum(p / p.sum(), 0)
    cols, labels = [], []
    t0 = time.time()
    for c in range(N // CHUNK):
        dense = torch.randint(0, 64, (CHUNK, 13), generator=g, device=dev)
        hashed = 64 + torch.searchsorted(cdf, torch.rand((CHUNK, 26), generator=g, device=dev, dtype=torch.float64)).clamp_(max=D - 65)
        cc, _ = torch.sort(torch.cat([dense, hashed], 1), dim=1)
        cols.append(cc.to(torch.int32).reshape(-1))
        labels.append((torch.rand(CHUNK, generator=g, device=dev) < 0.03).to(torch.float64))
    cols = torch.cat(cols)
    labels = torch.cat(labels)
    vals = torch.ones(cols.numel(), dtype=torch.float32, device=dev)
    row_ptr = torch.arange(N + 1, dtype=torch.int64, device=dev) * 39
    torch.cuda.synchronize()
    print("generated %d rows, %d entries in %.1f s" % (N, cols.numel(), time.time() - t0), flush=True)
    t0 = time.time()
    ds = pml.Dataset.from_device_pointers(ctx, N, D, row_ptr.data_ptr(), cols.data_ptr(), vals.data_ptr(), pml.abi.F32,
                                          labels.data_ptr(),
                                          storage=pml.STORE_F32 if os.environ.get("STORAGE") == "f32" else pml.STORE_AUTO)
    ctx.synchronize()
    info = ds.info()
    print("value bytes per entry: %d" % info.value_bytes)
    print("packed in %.2f s; device bytes %.2f GB; tiling: %d hot cols, %d tiles x %d rows; peak torch+lib memory %.1f GB" %
          (time.time() - t0, info.device_bytes / 1e9, info.tiled_hot_cols, info.tiled_tiles, info.tiled_tile_rows,
           (torch.cuda.mem_get_info()[1] - torch.cuda.mem_get_info()[0]) / 1e9), flush=True)
    del cols, vals, row_ptr
    torch.cuda.empty_cache()
    coef = torch.randn(D, generator=g, device=dev, dtype=torch.float64) * 0.05
    v = torch.randn(D, generator=g, device=dev, dtype=torch.float64)
    raw = torch.empty(D + 2, dtype=torch.float64, device=dev)
    raw2 = torch.empty(D + 2, dtype=torch.float64, device=dev)
    for _ in range(3):
        ds.value_and_gradient_raw_dev(pml.LOGISTIC, coef.data_ptr(), raw.data_ptr())
    ctx.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        ds.value_and_gradient_raw_dev(pml.LOGISTIC, coef.data_ptr(), raw.data_ptr())
    ctx.synchronize()
    per = (time.perf_counter() - t0) / 10
    print("value+gradient: %.3f ms per evaluation of the shard = %.1f evals/s per GPU; algorithmic %.2f GB -> %.0f GB/s" %
          (1e3 * per, 1 / per, info.vg_algorithmic_bytes / 1e9 - 2 * N * 8 / 1e9, (info.vg_algorithmic_bytes - 2 * N * 8) / per / 1e9),
          flush=True)
    ctx.set_profiling(True)
    ds.value_and_gradient_raw_dev(pml.LOGISTIC, coef.data_ptr(), raw.data_ptr())
    print({k: round(v_[0], 3) for k, v_ in ctx.read_profile().items() if v_[1]}, flush=True)
    ctx.set_profiling(False)
    # self-checks: determinism, Hessian-vector linearity, objective at zero = n log 2
    ds.value_and_gradient_raw_dev(pml.LOGISTIC, coef.data_ptr(), raw2.data_ptr())
    ctx.synchronize()
    assert torch.equal(raw, raw2)
    zero = torch.zeros(D, dtype=torch.float64, device=dev)
    ds.value_and_gradient_raw_dev(pml.LOGISTIC, zero.data_ptr(), raw.data_ptr())
    ctx.synchronize()
    assert abs(raw[D].item() - N * np.log(2.0)) <= 1e-9 * N
    h = [torch.empty(D + 2, dtype=torch.float64, device=dev) for _ in range(3)]
    ds.hessian_vector_raw_dev(pml.LOGISTIC, coef.data_ptr(), v.data_ptr(), h[0].data_ptr(), tag=1)
    ds.hessian_vector_raw_dev(pml.LOGISTIC, coef.data_ptr(), coef.data_ptr(), h[1].data_ptr(), tag=1)
    w = (2.0 * v - 0.5 * coef).contiguous()
    ds.hessian_vector_raw_dev(pml.LOGISTIC, coef.data_ptr(), w.data_ptr(), h[2].data_ptr(), tag=1)
    ctx.synchronize()
    err = (h[2][:D] - (2.0 * h[0][:D] - 0.5 * h[1][:D])).abs().max().item() / h[2][:D].abs().max().item()
    print("Hv linearity rel err %.2e" % err)
    assert err <= 1e-11
    print("ok")
    if os.environ.get("OWLQN"):
        # configs[2]'s optimizer on this shard alone: OWL-QN, L1 (alpha = 1) and elastic net (alpha = 0.5), lambda = 1,
        # optimizer state in device memory (the Breeze-style driver of photon_ml_b200/optimization.py)
        from photon_ml_b200.function import DeviceGLMLossFunction, NoNormalization
        from photon_ml_b200.optimization import OWLQN
        del h, w, raw, raw2
        for alpha in (1.0, 0.5):
            obj = DeviceGLMLossFunction(ctx, pml.LOGISTIC, l2RegWeight=(1.0 - alpha) or None)
            opt = OWLQN(alpha, NoNormalization(), isTrackingState=False)
            with obj.stream():
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                w_opt, f_opt = opt.optimize(obj, torch.zeros(D, dtype=torch.float64, device=dev), ds)
                torch.cuda.synchronize()
                wall = time.perf_counter() - t0
            print("OWLQN alpha=%.1f lambda=1: %.2f s, %d iterations, %d evaluations, %s, objective %.6f, %d of %d coefficients non-zero"
                  % (alpha, wall, opt.currentState.iter, opt.evaluations, opt.convergenceReason, f_opt,
                     int((w_opt != 0).sum()), D), flush=True)


if __name__ == "__main__":
    main()
