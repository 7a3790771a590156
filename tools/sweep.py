"""Experiment helper (not part of the product): times the evaluation kernels of the bench workload
for several row-pass variants, for the column-pass geometry given by $PML_COL_ITEMS."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
import photon_ml_b200 as pml  # noqa: E402


def main():
    rows = int(os.environ.get("SWEEP_ROWS", bench.N_ROWS))
    device = torch.device("cuda", 0)
    ctx = pml.Context(0)
    chunks = range(rows // bench.GEN_CHUNK)
    row_ptr, cols, vals, labels = bench.generate_rows(torch, device, chunks)
    torch.cuda.synchronize()
    t0 = time.time()
    ds = pml.Dataset.from_device_pointers(ctx, labels.numel(), bench.N_FEATS, row_ptr.data_ptr(), cols.data_ptr(),
                                          vals.data_ptr(), pml.abi.F32, labels.data_ptr(), storage=pml.STORE_F32)
    ctx.synchronize()
    print("pack %.3fs items=%s" % (time.time() - t0, os.environ.get("PML_COL_ITEMS", "16")), flush=True)
    del row_ptr, cols, vals, labels
    info = ds.info()
    d = info.d
    coef = torch.randn(d, device=device, dtype=torch.float64) * 0.02
    v = torch.randn(d, device=device, dtype=torch.float64)
    raw = torch.empty(d + 2, dtype=torch.float64, device=device)
    torch.cuda.synchronize()
    rb, cb = bench.kernel_bytes(info, 0, 0)
    variants = [int(x) for x in os.environ.get("SWEEP_VARIANTS", "1,8").split(",")]
    for var in variants:
        ds.set_margin_variant(var)
        for _ in range(3):
            ds.value_and_gradient_raw_dev(1, coef.data_ptr(), raw.data_ptr())
        ctx.read_profile()
        ctx.set_profiling(True)
        for _ in range(10):
            ds.value_and_gradient_raw_dev(1, coef.data_ptr(), raw.data_ptr())
        prof = ctx.read_profile()
        for _ in range(5):
            ds.hessian_vector_raw_dev(1, coef.data_ptr(), v.data_ptr(), raw.data_ptr())
        hv = ctx.read_profile()
        ctx.set_profiling(False)
        r = prof["row_pass"][0] / prof["row_pass"][1]
        c = prof["col_pass"][0] / prof["col_pass"][1]
        p = prof["prepare_vectors"][0] / max(1, prof["prepare_vectors"][1])
        h = hv["row_pass"][0] / hv["row_pass"][1]
        print("variant %2d: row %.3f ms (%.0f GB/s)  col %.3f ms (%.0f GB/s)  prepare %.3f ms  hv-row(cold) %.3f ms" %
              (var, r, rb / r / 1e6, c, cb / c / 1e6, p, h), flush=True)


if __name__ == "__main__":
    main()
