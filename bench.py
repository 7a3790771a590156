#!/usr/bin/env python
"""bench.py -- value+gradient evaluations/s of the GLM objective path on B200.

Workload (BASELINE.json configs[1], the configuration the metric is quoted on): Poisson
regression, synthetic 10 M rows x 1 M features, 66 Zipf(1.1) draws per row (duplicates merged,
so ~50 stored entries per row), values in {1,2,3,4}/4 (fp32-exact), offsets 0, weights 1.  A
"step" is one value+gradient evaluation over all rows.  With --gpus N the SAME rows are sharded
over N ranks (strong scaling); every rank produces raw partial sums and one merge of d+2 doubles
(the library's peer-memory kernel; PML_BENCH_COLLECTIVE=nccl for torch's NCCL all-reduce) replaces
Spark's treeAggregate.

  python bench.py [--gpus N] [--steps K] [--warmup W]                 our arm
  python bench.py --impl reference [--gpus N] [--steps K] [--warmup W]  CPU restatement of the
      reference loop on the host cores (Spark local[N] shape; no JVM on this box)

One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_ROWS = 10_000_000
N_FEATS = 1_000_000
DRAWS_PER_ROW = 66          # Zipf(1.1) duplicates merge: 66 draws leave ~50 stored entries per row
ZIPF_A = 1.1
GEN_CHUNK = 250_000          # rows per generator chunk; 40 chunks split evenly over 1/2/4/8 ranks
SEED = 2
LOSS_POISSON = 1


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def generate_chunk(torch, device, chunk_id, cdf, w_true):
    """One generator chunk: (lens[int64 rows], cols[int32], vals[float32], labels[float64])."""
    g = torch.Generator(device=device)
    g.manual_seed(SEED * 1_000_003 + chunk_id)
    rows = GEN_CHUNK
    u = torch.rand((rows, DRAWS_PER_ROW), generator=g, device=device, dtype=torch.float64)
    cols = torch.searchsorted(cdf, u).clamp_(max=N_FEATS - 1)
    cols, _ = torch.sort(cols, dim=1)
    keep = torch.ones_like(cols, dtype=torch.bool)
    keep[:, 1:] = cols[:, 1:] != cols[:, :-1]
    vals = torch.randint(1, 5, (rows, DRAWS_PER_ROW), generator=g, device=device).to(torch.float64) / 4.0
    margin = (vals * w_true[cols] * keep).sum(dim=1).clamp_(-5.0, 5.0)
    labels = torch.poisson(torch.exp(margin), generator=g)
    lens = keep.sum(dim=1)
    return lens, cols[keep].to(torch.int32), vals[keep].to(torch.float32), labels


def zipf_cdf():
    """The Zipf(1.1) CDF over the columns, built ONCE on the host with numpy: a prefix sum on the GPU is not
    reproducible from launch to launch (round 1: one flipped searchsorted boundary moved the objective of an
    8-GPU run by 8e-11), so every rank uploads the same bits instead of computing its own."""
    p = np.arange(1, N_FEATS + 1, dtype=np.float64) ** -ZIPF_A
    return np.cumsum(p / p.sum())


def data_checksum(torch, row_ptr, cols, vals, labels):
    """Exact in any summation order (integers and multiples of 1/4 far below 2^53): equal totals over all ranks
    for every N mean every N evaluated the same rows."""
    return [int(labels.numel()), int(cols.numel()), int(cols.sum(dtype=torch.int64)), float(vals.sum(dtype=torch.float64)),
            float(labels.sum())]


def generate_rows(torch, device, chunk_ids):
    cdf = torch.from_numpy(zipf_cdf()).to(device)
    g = torch.Generator(device=device)
    g.manual_seed(SEED)
    w_true = torch.randn(N_FEATS, generator=g, device=device, dtype=torch.float64) * 0.1
    w_true = w_true * (torch.rand(N_FEATS, generator=g, device=device) < 0.01)
    lens, cols, vals, labels = [], [], [], []
    for c in chunk_ids:
        a, b, v, y = generate_chunk(torch, device, c, cdf, w_true)
        lens.append(a); cols.append(b); vals.append(v); labels.append(y)
    lens = torch.cat(lens)
    row_ptr = torch.zeros(lens.numel() + 1, dtype=torch.int64, device=device)
    torch.cumsum(lens, dim=0, out=row_ptr[1:])
    return row_ptr, torch.cat(cols), torch.cat(vals), torch.cat(labels)


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu_index = gpu_index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu_index), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, flag in zip(names, f[5:9]):
                if flag.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


NCU_FULL_CAPTURES = ("r02_sell_ncu_full.csv", "r02_ncu_full.csv", "r01_final3_ncu_full.csv", "r01_tiled_ncu_full.csv")   # newest first


def ncu_traffic(kernel_substrings):
    """DRAM bytes (read + write) per launch of the named kernels, summed, from the committed `ncu --set full`
    captures of this same workload (profiles/, the newest capture that holds the kernel); (None, []) when one of them
    is in no capture."""
    import csv
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total, found, used = 0.0, set(), []
    for name in NCU_FULL_CAPTURES:
        path = os.path.join(ROOT, "profiles", name)
        if not os.path.exists(path):
            continue
        rows = list(csv.reader(open(path)))
        hdr, units = rows[0], rows[1]
        ki, ri, wi = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        for r in rows[2:]:
            for sub in kernel_substrings:
                if sub in r[ki] and sub not in found:
                    total += float(r[ri]) * scale[units[ri]] + float(r[wi]) * scale[units[wi]]
                    found.add(sub)
                    if name not in used:
                        used.append(name)
    return (total, used) if len(found) == len(kernel_substrings) else (None, [])


def kernel_bytes(info, has_offset, has_weight):
    """ALGORITHMIC bytes per launch of the two evaluation kernels."""
    V = info.value_bytes
    n, d, nnz, m = info.n, info.d, info.nnz, info.nonempty_cols
    row = nnz * (V + 4) + (n + 1) * 4 + n * 8 * (1 + has_offset + has_weight) + n * 8 + d * 8
    col = nnz * (V + 4) + n * 8 + m * 8 + (info.n_chunks + 1) * 4 + d * 8
    return row, col


def rel_inf(a, b):
    """max |a - b| / max |b|"""
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def pack_rows(pml, ctx, torch, rows):
    row_ptr, cols, vals, labels = rows
    ds = pml.Dataset.from_device_pointers(ctx, labels.numel(), N_FEATS, row_ptr.data_ptr(), cols.data_ptr(), vals.data_ptr(),
                                          pml.abi.F32, labels.data_ptr(), storage=pml.STORE_F32)
    ctx.synchronize()
    return ds


def rows_to_host(rows):
    """Device CSR -> the oracle's host arrays (values widened to the Doubles the reference holds)."""
    row_ptr, cols, vals, labels = rows
    return (row_ptr.cpu().numpy(), cols.cpu().numpy(), vals.cpu().numpy().astype(np.float64), labels.cpu().numpy())


def run_ours(args):
    import torch
    import torch.distributed as dist
    import photon_ml_b200 as pml

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node %d for --gpus %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    ctx = pml.Context(local_rank)
    ext = torch.cuda.ExternalStream(ctx.stream, device=device)

    n_chunks = N_ROWS // GEN_CHUNK
    mine = list(range(rank * n_chunks // world, (rank + 1) * n_chunks // world))
    t0 = time.time()
    rows = generate_rows(torch, device, mine)
    torch.cuda.synchronize()
    t_gen = time.time() - t0
    checksum = torch.tensor(data_checksum(torch, *rows), dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(checksum)
    checksum = checksum.tolist()
    t0 = time.time()
    ds = pack_rows(pml, ctx, torch, rows)
    t_pack = time.time() - t0
    host_rows = None
    if world == 1 and rank == 0 and not (args.no_cpu_baseline and args.no_parity):
        host_rows = rows_to_host(rows)            # the same rows for the oracle (parity) and the CPU baseline
    del rows
    torch.cuda.empty_cache()
    info = ds.info()
    d = info.d

    gen = torch.Generator(device=device)
    gen.manual_seed(99)
    coef = torch.randn(d, generator=gen, device=device, dtype=torch.float64) * 0.02
    raw = torch.empty(d + 2, dtype=torch.float64, device=device)
    out_value = torch.empty(1, dtype=torch.float64, device=device)
    out_grad = torch.empty(d, dtype=torch.float64, device=device)
    torch.cuda.synchronize()

    # The exchange step that replaces treeAggregate (ValueAndGradientAggregator.scala:243-247): by default the
    # library's own merge kernel over NVLink peer memory, fused with getVector (peer.cu); PML_BENCH_COLLECTIVE=nccl
    # (or a box without peer access) takes torch's NCCL all-reduce of the d+2 raw sums followed by finalize_dev.
    peer, collective = None, "none"
    if world > 1:
        collective = "nccl all-reduce of d+2 doubles, then finalize_dev"
        if os.environ.get("PML_BENCH_COLLECTIVE", "peer") == "peer":
            try:
                peer = pml.Peer(ctx, d)
            except Exception as e:  # noqa: BLE001
                sys.stderr.write("rank %d: peer memory unavailable (%s)\n" % (rank, e))
            ok = torch.tensor([1 if peer is not None else 0], device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok) == 0:
                if peer is not None:
                    peer.close()
                peer = None
            else:
                if peer.transport == "multicast":
                    collective = ("peer_merge_mc_kernel: in-switch sums (multimem.ld_reduce over an NVSwitch multicast "
                                  "object) fused with getVector, results fanned out by multimem.st")
                else:
                    collective = ("peer_merge_kernel: sums over NVLink peer memory in rank order, fused with getVector"
                                  " (multicast unavailable: %s)" % getattr(peer, "multicast_error", "disabled"))
                result = torch.as_tensor(peer.result_array(), device=device)
                out_grad, out_value = result[:d], result[d:d + 1]   # the library's result buffer, no copy

    def step():
        if peer is not None:
            peer.value_and_gradient(ds, LOSS_POISSON, coef.data_ptr())
            return
        ds.value_and_gradient_raw_dev(LOSS_POISSON, coef.data_ptr(), raw.data_ptr())
        if world > 1:
            dist.all_reduce(raw)
        ds.finalize_dev(LOSS_POISSON, raw.data_ptr(), out_value.data_ptr(), out_grad.data_ptr())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(ext):
        for _ in range(args.warmup):
            step()
        barrier()
        ctx.read_profile()
        # The timed region replays the captured CUDA graph of the evaluation -- what every caller of the library
        # gets -- at every N, so the 1 -> 8 curve uses one launch mode.  The per-kernel CUDA events need plain
        # launches: they are recorded over a second loop of the same K steps right after (live, same process).
        ctx.set_profiling(False)
        sampler = ClockSampler(local_rank)
        if rank == 0:
            sampler.start()
        launches0 = ctx.launch_count
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        start.record(ext)
        for _ in range(args.steps):
            step()
        stop.record(ext)
        barrier()
        ms = start.elapsed_time(stop)
        launches = ctx.launch_count - launches0
        ctx.set_profiling(True)
        pa, pb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        pa.record(ext)
        for _ in range(args.steps):
            step()
        pb.record(ext)
        barrier()
        profiled_ms = pa.elapsed_time(pb) / args.steps
        clocks = sampler.stop() if rank == 0 else None
        profile = ctx.read_profile()
        ctx.set_profiling(False)

        # ---- end to end through the host-buffer C ABI: pinned host coefficients in, gradient out
        host_peer = peer is not None and hasattr(peer, "value_and_gradient_host")
        shared = None
        if host_peer:
            # One host buffer for all ranks (the driver's coefficient / gradient arrays, as a file in /dev/shm mapped
            # and page-locked by every rank): rank 0 owns the vectors, every rank moves its slice, and a flag word per
            # rank tells the driver when the last slice has landed.
            shared = SharedHostVectors(torch, dist, rank, world, d, coef.cpu().numpy())
            h_coef_np, h_grad_np = shared.coef, shared.grad
        else:
            h_coef = torch.empty(d, dtype=torch.float64).pin_memory()
            h_coef.copy_(coef.cpu())
            h_grad = torch.empty(d, dtype=torch.float64).pin_memory()
            h_raw = torch.empty(d + 2, dtype=torch.float64).pin_memory()
            h_coef_np, h_grad_np = h_coef.numpy(), h_grad.numpy()
        e2e_iter = [0]

        def e2e_step():
            if world == 1:
                return ds.value_and_gradient(LOSS_POISSON, h_coef_np, out=h_grad_np)[0]
            if host_peer:
                # every rank moves 1/N of the coefficients up and 1/N of the gradient down over its own PCIe link;
                # the slices travel between the GPUs over NVLink inside the call
                value = peer.value_and_gradient_host(ds, LOSS_POISSON, h_coef_np, h_grad_np)
                e2e_iter[0] += 1
                shared.arrive_and_wait(e2e_iter[0])
                return value
            # the driver (rank 0) owns the host buffers: one upload, an NVLink broadcast of the coefficients
            # (what sc.broadcast does in the reference, DistributedObjectiveFunction.scala:57-58), the
            # evaluation with its all-reduce, and one download of gradient + value on rank 0
            if rank == 0:
                coef.copy_(h_coef, non_blocking=True)
            dist.broadcast(coef, 0)
            step()
            if rank == 0:
                h_grad.copy_(out_grad, non_blocking=True)
                h_raw[:1].copy_(out_value, non_blocking=True)
            ext.synchronize()
            return float(h_raw[0]) if rank == 0 else 0.0

        for _ in range(max(3, args.warmup)):
            e2e_step()
        barrier()
        e2e_steps = args.steps
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_value = e2e_step()
        barrier()
        e2e_s = time.perf_counter() - t0

        # Hessian-vector products (TRON's inner loop): cold and with the cached margin pass
        v = torch.randn(d, generator=gen, device=device, dtype=torch.float64)
        hv_ms = {}
        for name, tag in (("cold", 0), ("cached", 12345)):
            def hv_step():
                if peer is not None:
                    peer.hessian_vector(ds, LOSS_POISSON, coef.data_ptr(), v.data_ptr(), tag=tag)
                    return
                ds.hessian_vector_raw_dev(LOSS_POISSON, coef.data_ptr(), v.data_ptr(), raw.data_ptr(), tag=tag)
                if world > 1:
                    dist.all_reduce(raw)
                ds.finalize_dev(LOSS_POISSON, raw.data_ptr(), None, out_grad.data_ptr())
            for _ in range(3):
                hv_step()
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(ext)
            for _ in range(max(3, args.steps // 2)):
                hv_step()
            b.record(ext)
            barrier()
            hv_ms[name] = a.elapsed_time(b) / max(3, args.steps // 2)
        e2e_grad = np.array(h_grad_np, copy=True)
        if shared is not None:
            shared.close()
        hv_result = out_grad.clone()
        step()                      # leaves the value+gradient result in out_value / out_grad for the parity block
        barrier()

    # max over ranks
    times = torch.tensor([ms, e2e_s, profiled_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms, e2e_s, profiled_ms = (float(x) for x in times)

    parity = None
    if not args.no_parity:
        parity = parity_block(torch, dist, pml, ctx, ext, world, rank, device, ds, coef, v, out_value, out_grad, hv_result,
                              host_rows, e2e_value, e2e_grad)

    if rank == 0:
        peak, peak_src = peaks()
        row_bytes, col_bytes = kernel_bytes(info, 0, 0)
        per = lambda k: profile[k][0] / max(1, profile[k][1])   # noqa: E731  ms per launch
        # the CSR pass = the entry-chunked product + its row epilogue; the CSC pass = tiled hot columns + cold columns
        row_ms = per("row_pass") + per("row_epilogue")
        col_ms = per("col_pass") + per("col_pass_tiled")
        sliced_ell = info.sliced_ell_padding_permille >= 0
        row_kernels = ("CSR pass X w: sell_rows_kernel (one lane per row, sliced ELL) + row_epilogue_kernel" if sliced_ell else
                       "CSR pass X w: seg_spmv_warp_kernel + row_epilogue_kernel" if info.margin_variant == 1 else
                       "row_pass_kernel<%d lanes/row> (CSR pass, X w + pointwise)" % info.margin_variant)
        dominant = ("CSC pass X^T r: seg_spmv_tiled_kernel (hot columns) + seg_spmv_warp_kernel (cold columns)" if col_ms >= row_ms
                    else row_kernels)
        dom_bytes, dom_ms = (col_bytes, col_ms) if col_ms >= row_ms else (row_bytes, row_ms)
        traffic, traffic_files = None, []
        row_traffic = None
        if world == 1:
            row_traffic, row_files = ncu_traffic(["sell_rows_kernel" if sliced_ell else "0, 0, 1, 1024>", "row_epilogue_kernel"])
            traffic, traffic_files = (ncu_traffic(["seg_spmv_tiled_kernel", "0, 0, 0, 256>"]) if col_ms >= row_ms
                                      else (row_traffic, row_files))
        achieved = dom_bytes / (dom_ms * 1e-3) / 1e9
        ms_per_step = ms / args.steps
        step_bytes = info.vg_algorithmic_bytes - 2 * info.n * 8  # no offsets / weights in this workload
        floor_bytes = 8 * info.nnz + 28 * info.n + 16 * d        # one CSR pass, whatever the algorithm
        out = {
            "metric": "value+gradient evals/s",
            "value": 1e3 / ms_per_step,
            "unit": "evals/s",
            "n_gpus": world,
            "steps": args.steps,
            "warmup": args.warmup,
            "ms_per_step": ms_per_step,
            "higher_is_better": True,
            "scaling": "strong",
            "vs_baseline": None,
            "dtype": "f64",
            "data": "synthetic",
            "config": {
                "workload": "configs[1]: Poisson regression, 10M rows x 1M feats, 66 Zipf(1.1) draws/row "
                            "(%.1f stored nnz/row), fp32 values, fp64 accumulation" % (info.nnz / max(1, info.n)),
                "rows_total": N_ROWS, "rows_per_gpu": int(info.n), "nnz_per_gpu": int(info.nnz), "d": int(d),
                "parallelism": "row-sharded x%d, one merge of d+2 fp64 per evaluation" % world if world > 1 else "single GPU",
                "collective": collective,
                "l2_flush": "inputs larger than L2 (%.1f GB streamed per step vs 126 MB L2)" % (step_bytes / 1e9),
                "row_pass_variant": int(info.margin_variant),
                "launch_mode": "CUDA graph replay in the timed region (every N); per-kernel CUDA events over a second loop "
                               "of the same %d steps with plain launches (%.4f ms per step there)" % (args.steps, profiled_ms),
                "csc_tiling": {"hot_cols": int(info.tiled_hot_cols), "tiles": int(info.tiled_tiles),
                               "tile_rows": int(info.tiled_tile_rows)},
                "data_checksum": dict(zip(("rows", "nnz", "sum_col", "sum_val", "sum_label"), checksum)),
                "generate_s": round(t_gen, 2), "pack_s": round(t_pack, 2),
            },
            "e2e": {"value": e2e_steps / e2e_s, "unit": "evals/s",
                    "h2d_bytes_per_step": d * 8, "d2h_bytes_per_step": d * 8 + 8 * world,
                    "ms_per_step": 1e3 * e2e_s / e2e_steps,
                    "api": "pml_value_and_gradient (host buffers, pinned)" if world == 1 else
                           ("pml_value_and_gradient_allreduce (host buffers): every rank uploads 1/N of the coefficients and "
                            "downloads 1/N of the gradient over its own PCIe link, slices exchanged over NVLink"
                            if host_peer else
                            "rank 0: H2D coef; NCCL broadcast of coef; evaluation + merge (%s); rank 0: D2H" % collective),
                    "objective": e2e_value},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "frac_of_nominal_8000": achieved / 8000.0, "peak_source": peak_src,
                         "traffic": traffic, "traffic_source": "profiles/%s (ncu --set full, dram read+write)" % (
                             " + ".join(traffic_files) or "-"),
                         "algorithmic_bytes_per_launch": int(dom_bytes),
                         "avg_launch_ms": dom_ms},
            "kernels": {k: {"ms_per_launch": v[0] / max(1, v[1]), "launches": int(v[1])} for k, v in profile.items()},
            "row_pass": {"kernel": row_kernels, "algorithmic_bytes": int(row_bytes), "ms": row_ms,
                         "GBps": row_bytes / (row_ms * 1e-3) / 1e9, "frac": row_bytes / (row_ms * 1e-3) / 1e9 / peak,
                         "traffic": row_traffic, "sliced_ell_padding_permille": int(info.sliced_ell_padding_permille)},
            "col_pass": {"algorithmic_bytes": int(col_bytes), "ms": col_ms, "GBps": col_bytes / (col_ms * 1e-3) / 1e9,
                         "frac": col_bytes / (col_ms * 1e-3) / 1e9 / peak},
            "step": {"algorithmic_bytes_two_pass": int(step_bytes), "GBps": step_bytes / (ms_per_step * 1e-3) / 1e9,
                     "frac": step_bytes / (ms_per_step * 1e-3) / 1e9 / peak,
                     "one_pass_floor_bytes": int(floor_bytes),
                     "one_pass_floor_GBps": floor_bytes / (ms_per_step * 1e-3) / 1e9,
                     "one_pass_floor_frac": floor_bytes / (ms_per_step * 1e-3) / 1e9 / peak},
            "step_algorithmic_GBps": step_bytes / (ms_per_step * 1e-3) / 1e9,
            "hessian_vector": hessian_vector_report(info, d, hv_ms, peak),
            "parity": parity,
        }
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(host_rows, steps=20, warmup=5)
        if world == 1 and not args.no_convergence:
            out["convergence"] = convergence(ctx, ds, d, host_rows)
        host_rows = None
        if world == 1 and not args.no_convergence:
            out["random_effect_batch"] = random_effect_batch(ctx)
        emit(out)
    ds.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if parity is not None and not parity.get("ok", True):
        raise SystemExit("parity block failed: %s" % json.dumps(parity))


class SharedHostVectors:
    """Coefficients and gradient of the end-to-end step in ONE host buffer shared by the rank processes (a file in
    /dev/shm, page-locked by every rank with cudaHostRegister), plus a flag word per rank: rank r bumps flags[r] once
    its slice of the gradient has landed, and everybody waits for all flags -- the driver then holds the full vector."""

    def __init__(self, torch, dist, rank, world, d, coef):
        self.rank, self.world, self.torch = rank, world, torch
        self.path = "/dev/shm/pml_bench_%s_%s.bin" % (os.environ.get("MASTER_PORT", "0"), os.getppid() if world > 1 else os.getpid())
        words = 2 * d + 64
        if rank == 0:
            np.zeros(words).tofile(self.path)
        dist.barrier()
        self.map = np.memmap(self.path, dtype=np.float64, mode="r+", shape=(words,))
        self.coef, self.grad = self.map[:d], self.map[d:2 * d]
        self.flags = self.map[2 * d:2 * d + world]
        if rank == 0:
            self.coef[:] = coef
        self.registered = torch.cuda.cudart().cudaHostRegister(self.map.ctypes.data, self.map.nbytes, 0)
        dist.barrier()

    def arrive_and_wait(self, it):
        self.flags[self.rank] = it
        while float(self.flags.min()) < it:
            pass

    def close(self):
        try:
            self.torch.cuda.cudart().cudaHostUnregister(self.map.ctypes.data)
        except Exception:  # noqa: BLE001
            pass
        if self.rank == 0:
            try:
                os.unlink(self.path)
            except OSError:
                pass


def hessian_vector_report(info, d, hv_ms, peak):
    """Hessian-vector evaluations/s with their own algorithmic bytes (no offsets / weights here)."""
    n, nnz = info.n, info.nnz
    cold_bytes = 16 * nnz + 20 * n + 28 * d      # two sparse streams, r written and read, w, v, Hv
    cached_bytes = 16 * nnz + 28 * n + 20 * d    # the cached curvature D read instead of the margin work
    return {"cold_evals_per_s": 1e3 / hv_ms["cold"], "cached_evals_per_s": 1e3 / hv_ms["cached"],
            "cold_ms": hv_ms["cold"], "cached_ms": hv_ms["cached"],
            "roofline": {"bound": "hbm", "peak": peak, "unit": "GB/s",
                         "cold": {"algorithmic_bytes": int(cold_bytes), "achieved": cold_bytes / hv_ms["cold"] / 1e6,
                                  "frac": cold_bytes / hv_ms["cold"] / 1e6 / peak},
                         "cached": {"algorithmic_bytes": int(cached_bytes), "achieved": cached_bytes / hv_ms["cached"] / 1e6,
                                    "frac": cached_bytes / hv_ms["cached"] / 1e6 / peak}}}


PARITY_SAMPLE_ROWS = 500_000   # natural tiling geometry (tiling starts at 262 144 rows), small enough for the sequential oracle


def parity_block(torch, dist, pml, ctx, ext, world, rank, device, ds, coef, v, out_value, out_grad, hv_result, host_rows,
                 e2e_value, e2e_grad):
    """What this run's numbers are worth: the evaluated result against an independent evaluation of the same rows.
    N = 1: the GPU value / gradient over all 10 M rows against the oracle (the reference's row loop, partitioned) on
    the same rows, and value / gradient / Hessian-vector on the first 500 k rows -- packed on their own, so in the
    geometry the packer picks by itself -- against the sequential oracle.
    N > 1: the merged result against ONE GPU evaluating all the rows (rank 0 packs them, untimed), and bit-equality
    of the result across the ranks."""
    import zlib
    tol = 1e-12
    value = float(out_value.cpu()[0])
    grad = out_grad.cpu().numpy().copy()
    hv = hv_result.cpu().numpy()
    out = {"tolerance": tol}
    if world == 1:
        if host_rows is None:
            return None
        import oracle
        row_ptr, cols, vals, labels = host_rows
        coef_h, v_h = coef.cpu().numpy(), v.cpu().numpy()
        cores = oracle.max_threads()
        full = oracle.Data(N_ROWS, N_FEATS, row_ptr, cols, vals, labels)
        f_ref, g_ref = oracle.value_and_gradient(full, oracle.POISSON, coef_h, parts=cores)
        out["full_rows_vs_oracle"] = {"rows": N_ROWS, "oracle": "partitioned x%d" % cores, "value": value, "oracle_value": f_ref,
                                      "value_rel": abs(value - f_ref) / abs(f_ref), "gradient_rel_inf": rel_inf(grad, g_ref),
                                      "e2e_value_rel": abs(e2e_value - f_ref) / abs(f_ref),
                                      "e2e_gradient_rel_inf": rel_inf(e2e_grad, g_ref)}
        ns = PARITY_SAMPLE_ROWS
        end = int(row_ptr[ns])
        sample = oracle.Data(ns, N_FEATS, row_ptr[:ns + 1], cols[:end], vals[:end], labels[:ns])
        f_s, g_s = oracle.value_and_gradient(sample, oracle.POISSON, coef_h)
        hv_s = oracle.hessian_vector(sample, oracle.POISSON, coef_h, v_h)
        dev = [torch.from_numpy(np.ascontiguousarray(a)).to(device) for a in
               (row_ptr[:ns + 1], cols[:end], vals[:end].astype(np.float32), labels[:ns])]
        small = pack_rows(pml, ctx, torch, dev)
        si = small.info()
        f_g, g_g = small.value_and_gradient(pml.POISSON, coef_h)
        hv_g = small.hessian_vector(pml.POISSON, coef_h, v_h)
        hv_g2 = small.hessian_vector(pml.POISSON, coef_h, v_h, tag=7)
        hv_g3 = small.hessian_vector(pml.POISSON, coef_h, v_h, tag=7)     # from the cached curvature
        small.close()
        out["sample_vs_oracle"] = {"rows": ns, "oracle": "sequential",
                                   "csc_tiling": {"hot_cols": int(si.tiled_hot_cols), "tiles": int(si.tiled_tiles),
                                                  "tile_rows": int(si.tiled_tile_rows)},
                                   "value_rel": abs(f_g - f_s) / abs(f_s), "gradient_rel_inf": rel_inf(g_g, g_s),
                                   "hessian_vector_rel_inf": max(rel_inf(hv_g, hv_s), rel_inf(hv_g2, hv_s)),
                                   "hessian_vector_cached_rel_inf": rel_inf(hv_g3, hv_s)}
        worst = max(out["full_rows_vs_oracle"]["value_rel"], out["full_rows_vs_oracle"]["gradient_rel_inf"],
                    out["full_rows_vs_oracle"]["e2e_value_rel"], out["full_rows_vs_oracle"]["e2e_gradient_rel_inf"],
                    *[x for k, x in out["sample_vs_oracle"].items() if k.endswith("_rel") or k.endswith("_rel_inf")])
        out["worst"] = worst
        out["ok"] = bool(worst <= tol)
        return out
    # ---- N > 1
    crc = zlib.crc32(grad.tobytes()) ^ zlib.crc32(np.float64(value).tobytes())
    crcs = torch.zeros(world, dtype=torch.int64, device=device)
    crcs[rank] = crc
    dist.all_reduce(crcs)
    if rank != 0:
        return {"ok": True}
    n_chunks = N_ROWS // GEN_CHUNK
    with torch.cuda.stream(ext):
        rows = generate_rows(torch, device, range(n_chunks))
        whole = pack_rows(pml, ctx, torch, rows)
        del rows
        g1 = torch.empty(N_FEATS, dtype=torch.float64, device=device)
        f1 = torch.empty(1, dtype=torch.float64, device=device)
        h1 = torch.empty(N_FEATS, dtype=torch.float64, device=device)
        whole.value_and_gradient_dev(LOSS_POISSON, coef.data_ptr(), f1.data_ptr(), g1.data_ptr())
        whole.hessian_vector_dev(LOSS_POISSON, coef.data_ptr(), v.data_ptr(), h1.data_ptr())
        ctx.synchronize()
        f_one, g_one, h_one = float(f1.cpu()[0]), g1.cpu().numpy(), h1.cpu().numpy()
        whole.close()
    out["merged_vs_single_gpu"] = {"rows": N_ROWS, "value": value, "single_gpu_value": f_one,
                                   "value_rel": abs(value - f_one) / abs(f_one), "gradient_rel_inf": rel_inf(grad, g_one),
                                   "hessian_vector_rel_inf": rel_inf(hv, h_one),
                                   "e2e_value_rel": abs(e2e_value - f_one) / abs(f_one),
                                   # the gradient the driver (rank 0) finds in its host buffer after the end-to-end step
                                   "e2e_gradient_rel_inf": rel_inf(e2e_grad, g_one)}
    out["result_crc32_per_rank"] = [int(x) for x in crcs.tolist()]
    out["ranks_bit_identical"] = len(set(out["result_crc32_per_rank"])) == 1
    worst = max(x for k, x in out["merged_vs_single_gpu"].items() if k.endswith("_rel") or k.endswith("_rel_inf"))
    out["worst"] = worst
    out["ok"] = bool(worst <= tol and out["ranks_bit_identical"])
    return out


def convergence(ctx, ds, d, host_rows=None):
    """Metric (3): optimizer wall time to convergence, driven through the host-buffer C ABI.
    TRON on this run's configs[1] data (Hessian-vector path) and LBFGS on configs[0] (a1a-shaped: the
    first 1605 rows of the reference's a9a fixture), the latter also on the CPU oracle under the same
    driver so that iteration counts can be compared."""
    import oracle
    import photon_ml_b200 as pml
    from photon_ml_b200.function import DistributedGLMLossFunction, NoNormalization
    from photon_ml_b200.optimization import LBFGS, TRON
    from tests import util
    from tests.objectives import OracleGLMObjective
    import torch
    from photon_ml_b200.function import DeviceGLMLossFunction
    out = {}
    dev = torch.device("cuda", ctx.device)

    def tron_device(lam, max_iter):
        obj = DeviceGLMLossFunction(ctx, LOSS_POISSON, l2RegWeight=lam)
        tron = TRON(NoNormalization(), maxNumIterations=max_iter)
        with obj.stream():
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            _, f = tron.optimize(obj, torch.zeros(d, dtype=torch.float64, device=dev), ds)
            torch.cuda.synchronize()
            wall = time.perf_counter() - t0
        return {"wall_s": wall, "iterations": tron.currentState.iter, "reason": tron.convergenceReason,
                "value_gradient_evals": tron.evaluations, "hessian_vector_evals": tron.hessian_vector_evaluations,
                "cg_iterations": int(sum(tron.cg_iterations)), "objective": f}

    def tron_host(objective, data, dim, max_iter):
        tron = TRON(NoNormalization(), maxNumIterations=max_iter)
        t0 = time.perf_counter()
        _, f = tron.optimize(objective, np.zeros(dim), data)
        return {"wall_s": time.perf_counter() - t0, "iterations": tron.currentState.iter, "reason": tron.convergenceReason,
                "value_gradient_evals": tron.evaluations, "hessian_vector_evals": tron.hessian_vector_evaluations,
                "cg_iterations": int(sum(tron.cg_iterations)), "objective": f}

    # (a) the reference's defaults (TRON.scala:97-98: 15 iterations): on this problem that is the iteration cap
    out["tron_configs1_poisson_l2_1_device_state"] = tron_device(1.0, 15)
    # (b) the same problem run TO CONVERGENCE (cap lifted to 300; FunctionValuesConverged, Optimizer.scala:136-150).  The
    #     per-iteration decrease hovers around the stopping threshold (1e-5 of the value at zero) for dozens of
    #     iterations on this ill-conditioned problem, so where exactly the run stops moves with the last bits of the
    #     evaluations: 79 iterations with the entry-chunked margin pass, more with the one-lane-per-row pass.
    out["tron_configs1_poisson_l2_1_device_state_to_convergence"] = tron_device(1.0, 300)
    # (c) a better conditioned setting, where the optimizer state in device memory and the drop-in shape (host vectors
    #     through the host-buffer ABI every evaluation, numpy optimizer state) can be compared count by count
    out["tron_configs1_poisson_l2_1000_device_state"] = tron_device(1000.0, 100)
    out["tron_configs1_poisson_l2_1000"] = tron_host(DistributedGLMLossFunction.create(LOSS_POISSON, l2RegWeight=1000.0), ds, d, 100)
    # (c') the metric's other optimizer on the same data: L-BFGS (Breeze-style driver, state in device memory)
    for lam in (1.0, 1000.0):
        obj = DeviceGLMLossFunction(ctx, LOSS_POISSON, l2RegWeight=lam)
        lb = LBFGS(NoNormalization(), maxNumIterations=100)
        with obj.stream():
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            _, f = lb.optimize(obj, torch.zeros(d, dtype=torch.float64, device=dev), ds)
            torch.cuda.synchronize()
            out["lbfgs_configs1_poisson_l2_%g_device_state" % lam] = {
                "wall_s": time.perf_counter() - t0, "iterations": lb.currentState.iter, "reason": lb.convergenceReason,
                "value_gradient_evals": lb.evaluations, "objective": f}
    # (d) the same driver on the CUDA objective and on the CPU oracle objective, first 500 k rows of the workload:
    #     iteration / evaluation / CG counts must agree (north star: identical iteration counts)
    if host_rows is not None:
        rp, cc, vv, yy = host_rows
        ns = PARITY_SAMPLE_ROWS
        end = int(rp[ns])
        sample_ds = pml.Dataset.from_csr(ctx, ns, N_FEATS, rp[:ns + 1], cc[:end], vv[:end], yy[:ns])
        sample_ref = oracle.Data(ns, N_FEATS, rp[:ns + 1], cc[:end], vv[:end], yy[:ns])
        lam = 50.0
        on_gpu = tron_host(DistributedGLMLossFunction.create(LOSS_POISSON, l2RegWeight=lam), sample_ds, N_FEATS, 100)
        on_cpu = tron_host(OracleGLMObjective(oracle.POISSON, l2RegWeight=lam), sample_ref, N_FEATS, 100)
        sample_ds.close()
        keys = ("iterations", "reason", "value_gradient_evals", "hessian_vector_evals", "cg_iterations")
        out["tron_first_500k_rows_poisson_l2_50_gpu_vs_cpu_oracle"] = {
            "gpu": on_gpu, "cpu_oracle": on_cpu, "same_counts": all(on_gpu[k] == on_cpu[k] for k in keys),
            "objective_rel_diff": abs(on_gpu["objective"] - on_cpu["objective"]) / abs(on_cpu["objective"])}
    n1, d1, row_ptr, col, val, label = util.read_libsvm(os.path.join(util.GOLDEN, "a9a_1605.libsvm"), d=123,
                                                        add_intercept=True)
    small = pml.Dataset.from_csr(ctx, n1, d1, row_ptr, col, val, label)
    ref = oracle.Data(n1, d1, row_ptr, col, val, label)
    runs = []
    for lam in (100.0, 10.0, 1.0, 0.1):   # README.md:252 regularization weights
        g, c = LBFGS(NoNormalization(), maxNumIterations=50), LBFGS(NoNormalization(), maxNumIterations=50)
        t0 = time.perf_counter()
        wg, fg = g.optimize(DistributedGLMLossFunction.create(pml.LOGISTIC, l2RegWeight=lam), np.zeros(d1), small)
        tg = time.perf_counter() - t0
        t0 = time.perf_counter()
        wc, fc = c.optimize(OracleGLMObjective(oracle.LOGISTIC, l2RegWeight=lam), np.zeros(d1), ref)
        tc = time.perf_counter() - t0
        runs.append({"lambda": lam, "gpu_wall_s": tg, "cpu_oracle_wall_s": tc, "gpu_iterations": g.currentState.iter,
                     "cpu_iterations": c.currentState.iter, "gpu_evals": g.evaluations, "cpu_evals": c.evaluations,
                     "us_per_eval_gpu": 1e6 * tg / g.evaluations,
                     "max_coef_diff": float(np.max(np.abs(wg - wc))), "objective": fg})
    out["lbfgs_configs0_a1a_logistic_l2"] = runs
    # the evaluation call alone (the Python optimizer loop above costs ~100 us per iteration by itself)
    w = np.random.default_rng(1).normal(scale=0.1, size=d1)
    g = np.empty(d1)
    lat = {}
    for _ in range(20):
        small.value_and_gradient(pml.LOGISTIC, w, out=g)
    t0 = time.perf_counter()
    for _ in range(300):
        small.value_and_gradient(pml.LOGISTIC, w, out=g)
    lat["gpu_host_buffer_call_us"] = 1e6 * (time.perf_counter() - t0) / 300
    t0 = time.perf_counter()
    for _ in range(300):
        oracle.value_and_gradient(ref, oracle.LOGISTIC, w)
    lat["cpu_oracle_1_thread_us"] = 1e6 * (time.perf_counter() - t0) / 300
    out["configs0_a1a_us_per_evaluation"] = lat
    small.close()
    return out


def make_entity_problems(rng, n_ent, k=10, d_orig=200, rows=None, device=None):
    """configs[4]-shaped per-entity problems: power-law rows per entity, k indicator features per row drawn
    from a d_orig-dimensional member-side shard, ~10 % positives -- and then what GAME does before training a random
    effect: every entity is projected onto the features ITS rows use (IndexMapProjectorRDD.scala:166-206; the same
    numbering as photon_ml_b200.io.IndexMapProjection, done with torch.unique on the GPU here because of the size), so
    per-entity dimensions come out at 10 ... ~40 and no projected column is empty."""
    import torch
    if rows is None:   # Pareto tail, mean ~3.3, capped ("power-law, mean 4")
        rows = np.minimum(np.floor(rng.random(n_ent) ** (-1.0 / 1.35)), 5000).astype(np.int64)
    ent_row_ptr = np.concatenate([[0], np.cumsum(rows)]).astype(np.int64)
    n = int(ent_row_ptr[-1])
    dev = device if device is not None else torch.device("cpu")
    gen = torch.Generator(device=dev)
    gen.manual_seed(int(rng.integers(1 << 31)))
    col = torch.sort(torch.randint(0, d_orig, (n, k), generator=gen, device=dev, dtype=torch.int64), dim=1)[0].reshape(-1)
    ent_of_entry = torch.repeat_interleave(torch.arange(n_ent, device=dev), torch.from_numpy(rows).to(dev) * k)
    uniq, inverse = torch.unique(ent_of_entry * d_orig + col, return_inverse=True)
    del col
    ent_dim = torch.bincount(uniq // d_orig, minlength=n_ent)
    coef_ptr = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), torch.cumsum(ent_dim, 0)])
    col_local = (inverse - coef_ptr[ent_of_entry]).to(torch.int32).cpu().numpy()
    del uniq, inverse, ent_of_entry
    row_ptr = np.arange(n + 1, dtype=np.int64) * k
    val = np.ones(n * k, dtype=np.float32)
    label = (rng.random(n) < 0.1).astype(np.float64)
    return ent_row_ptr, ent_dim.to(torch.int32).cpu().numpy(), row_ptr, col_local, val, label


def batch_bytes(n_ent, n, nnz, d_total, value_bytes):
    """ALGORITHMIC bytes of one batched value+gradient evaluation: the two entry streams, labels, the per-row
    multiplier and the per-row loss (each written once and read once), row / column / entity pointers, coefficients in,
    gradients and values out."""
    return 2 * nnz * (value_bytes + 4) + n * (8 + 16 + 16) + (n + 1) * 4 + d_total * 4 + d_total * 16 + n_ent * (8 + 8 + 8)


def random_effect_batch(ctx, n_ent=1_000_000, seed=5, solver_runs=("skip", "stream-all", "unfused-cg", "lbfgs")):
    """configs[4]-shaped per-entity batch (default: 1/10 scale, 1 M entities): every entity's value+gradient in one
    call, with its roofline, and every entity's L2-regularised problem solved in lock step."""
    import photon_ml_b200 as pml
    import torch
    rng = np.random.default_rng(seed)
    k = 10
    ent_row_ptr, ent_dim, row_ptr, col, val, label = make_entity_problems(rng, n_ent, k, device=torch.device("cuda", ctx.device))
    n = int(ent_row_ptr[-1])
    t0 = time.perf_counter()
    batch = pml.Batch(ctx, ent_row_ptr, ent_dim, row_ptr, col, val, label)
    t_pack = time.perf_counter() - t0
    del col, val
    coefs = rng.normal(scale=0.2, size=batch.d_total)
    for _ in range(3):
        batch.value_and_gradient(pml.LOGISTIC, coefs)
    reps = 10
    t0 = time.perf_counter()
    for _ in range(reps):
        batch.value_and_gradient(pml.LOGISTIC, coefs)
    per = (time.perf_counter() - t0) / reps
    dev = torch.device("cuda", ctx.device)
    t_coef = torch.tensor(coefs, device=dev)
    t_val = torch.empty(n_ent, dtype=torch.float64, device=dev)
    t_grad = torch.empty(batch.d_total, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    for _ in range(3):
        batch.value_and_gradient_dev(pml.LOGISTIC, t_coef.data_ptr(), t_val.data_ptr(), t_grad.data_ptr())
    ctx.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(ext)
    for _ in range(reps):
        batch.value_and_gradient_dev(pml.LOGISTIC, t_coef.data_ptr(), t_val.data_ptr(), t_grad.data_ptr())
    b.record(ext)
    ctx.synchronize()
    per_dev = a.elapsed_time(b) / reps * 1e-3
    ctx.read_profile()
    ctx.set_profiling(True)
    batch.value_and_gradient_dev(pml.LOGISTIC, t_coef.data_ptr(), t_val.data_ptr(), t_grad.data_ptr())
    ctx.synchronize()
    prof = ctx.read_profile()
    ctx.set_profiling(False)
    info = batch.dataset.info()
    peak, _ = peaks()
    nbytes = batch_bytes(n_ent, n, n * k, batch.d_total, info.value_bytes)
    out = {"entities": n_ent, "rows": n, "entries": n * k, "coefficients": int(batch.d_total), "pack_s": round(t_pack, 2),
           "ms_per_batch_evaluation_host_buffers": 1e3 * per, "entity_evaluations_per_s_host_buffers": n_ent / per,
           "ms_per_batch_evaluation_device_buffers": 1e3 * per_dev, "entity_evaluations_per_s_device_buffers": n_ent / per_dev,
           "device_ms": {kk: vv[0] for kk, vv in prof.items() if vv[1]},
           "roofline": {"bound": "hbm", "kernel": "pml_batch_value_and_gradient_dev (all kernels of one batched evaluation)",
                        "algorithmic_bytes": int(nbytes), "value_bytes_per_entry": int(info.value_bytes),
                        "achieved": nbytes / per_dev / 1e9, "peak": peak, "unit": "GB/s", "frac": nbytes / per_dev / 1e9 / peak}}
    # every entity's L2-regularised logistic problem solved to TRON's tolerance in lock step, state in device memory
    # (what RandomEffectCoordinate.updateModel does one entity at a time, RandomEffectCoordinate.scala:108-119);
    # "skip": converged entities are masked out and their rows are no longer streamed; "stream-all": as in round 1
    from photon_ml_b200.batched import BatchedTRON, REASONS
    del t_coef, t_val, t_grad
    # (one untimed solve first: the first Hessian-vector products allocate the curvature cache and capture the graphs)
    BatchedTRON(batch, pml.LOGISTIC, l2RegWeight=1.0, maxNumIterations=2).optimize()
    for mode in solver_runs:
        if mode == "lbfgs":
            continue
        solver = BatchedTRON(batch, pml.LOGISTIC, l2RegWeight=1.0, skip_converged=(mode != "stream-all"),
                             fused_cg=(mode != "unfused-cg"))
        ctx.synchronize()
        t0 = time.perf_counter()
        _, _, sinfo = solver.optimize()
        wall = time.perf_counter() - t0
        out["lockstep_tron_logistic_l2_1" + {"skip": "", "stream-all": "_streaming_converged_entities",
                                             "unfused-cg": "_cg_step_as_dots_and_axpys"}[mode]] = {
            "wall_s": wall, "entities_solved_per_s": n_ent / wall, "batched_value_gradient_evals": sinfo["evaluations"],
            "batched_hessian_vector_evals": sinfo["hessian_vector_evaluations"],
            "mean_iterations_per_entity": float(sinfo["iterations"].mean()),
            "reasons": {REASONS[c]: int((sinfo["reasons"] == c).sum()) for c in range(len(REASONS))}}
    # the same problems under the optimizer GAME's coordinates default to (OptimizerFactory.scala:46-62): lock-step L-BFGS
    if "lbfgs" in solver_runs:
        from photon_ml_b200.batched_lbfgs import BatchedLBFGS
        BatchedLBFGS(batch, pml.LOGISTIC, l2RegWeight=1.0, maxNumIterations=2).optimize()
        solver = BatchedLBFGS(batch, pml.LOGISTIC, l2RegWeight=1.0)
        ctx.synchronize()
        t0 = time.perf_counter()
        _, _, sinfo = solver.optimize()
        wall = time.perf_counter() - t0
        out["lockstep_lbfgs_logistic_l2_1"] = {
            "wall_s": wall, "entities_solved_per_s": n_ent / wall, "batched_value_gradient_evals": sinfo["evaluations"],
            "mean_iterations_per_entity": float(sinfo["iterations"].mean()),
            "mean_evaluations_per_entity": float(sinfo["entity_evaluations"].mean()),
            "reasons": {REASONS[c]: int((sinfo["reasons"] == c).sum()) for c in range(len(REASONS))}}
    batch.close()
    return out


def run_config4(args):
    """configs[4], the random-effect half: 10 M per-member problems sharded over the GPUs by entity with the
    reference's partitioner rule (bin packing of the heaviest entities, hash for the rest) and NO collective on the
    data path; every rank solves its entities in lock step.  value = entities solved per second, whole job."""
    import torch
    import torch.distributed as dist
    import photon_ml_b200 as pml
    from photon_ml_b200 import parallel
    from photon_ml_b200.batched import BatchedTRON, REASONS
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    n_total = int(os.environ.get("PML_BENCH_ENTITIES", "10000000"))
    # rows per member: Pareto tail with mean ~3.3, capped ("power-law, mean 4"); the same on every rank
    rows_all = np.minimum(np.floor(np.random.default_rng(5).random(n_total) ** (-1.0 / 1.35)), 5000).astype(np.int64)
    assignment = parallel.shard_entities(rows_all, world, capacity=10000)
    mine = np.flatnonzero(assignment == rank)
    ctx = pml.Context(local_rank)
    rng = np.random.default_rng(500 + rank)
    ent_row_ptr, ent_dim, row_ptr, col, val, label = make_entity_problems(rng, len(mine), rows=rows_all[mine], device=device)
    n = int(ent_row_ptr[-1])
    t0 = time.perf_counter()
    batch = pml.Batch(ctx, ent_row_ptr, ent_dim, row_ptr, col, val, label)
    t_pack = time.perf_counter() - t0
    del col, val
    ext = torch.cuda.ExternalStream(ctx.stream, device=device)
    t_coef = torch.randn(batch.d_total, dtype=torch.float64, device=device) * 0.2
    t_val = torch.empty(len(mine), dtype=torch.float64, device=device)
    t_grad = torch.empty(batch.d_total, dtype=torch.float64, device=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(ext):
        for _ in range(args.warmup):
            batch.value_and_gradient_dev(pml.LOGISTIC, t_coef.data_ptr(), t_val.data_ptr(), t_grad.data_ptr())
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        a.record(ext)
        for _ in range(args.steps):
            batch.value_and_gradient_dev(pml.LOGISTIC, t_coef.data_ptr(), t_val.data_ptr(), t_grad.data_ptr())
        b.record(ext)
        barrier()
        eval_ms = a.elapsed_time(b) / args.steps
    del t_coef, t_val, t_grad
    BatchedTRON(batch, pml.LOGISTIC, l2RegWeight=1.0, maxNumIterations=2).optimize()   # untimed: lazy buffers, graph captures
    solver = BatchedTRON(batch, pml.LOGISTIC, l2RegWeight=1.0)
    barrier()
    t0 = time.perf_counter()
    _, _, sinfo = solver.optimize()
    ctx.synchronize()
    wall = time.perf_counter() - t0
    # ---- the whole configuration: fixed effect (25 indicator features per row over d = 2 M, Zipf column popularity: 10^9
    # entries over all ranks) + the per-member random effect, coordinate descent fixed -> per-member.  A rank owns the
    # rows of ITS members for both coordinates, so the scores travel between the coordinates without leaving the GPU
    # (CoordinateDescent.scala:189,218 joins them by row id); only the fixed effect's sums cross ranks, through the
    # exchange over peer memory, and every rank trains the same fixed-effect model on the same merged numbers.
    cd = None
    if not args.no_convergence:
        from photon_ml_b200.game import CoordinateDescent, FixedEffect, RandomEffect
        D_FE, K_FE = 2_000_000, 25
        pz = np.arange(1, D_FE + 1, dtype=np.float64) ** -1.05
        cdf = torch.from_numpy(np.cumsum(pz / pz.sum())).to(device)
        gen = torch.Generator(device=device)
        gen.manual_seed(7_000 + rank)
        fe_cols = torch.empty((n, K_FE), dtype=torch.int32, device=device)
        for lo in range(0, n, 4_000_000):
            hi = min(n, lo + 4_000_000)
            c = torch.searchsorted(cdf, torch.rand((hi - lo, K_FE), generator=gen, device=device, dtype=torch.float64))
            fe_cols[lo:hi] = torch.sort(c.clamp_(max=D_FE - 1), dim=1)[0].to(torch.int32)
        fe_ptr = torch.arange(n + 1, dtype=torch.int64, device=device) * K_FE
        ones = torch.ones(n * K_FE, dtype=torch.float32, device=device)
        label_dev = torch.from_numpy(label).to(device)
        t0 = time.perf_counter()
        fe_ds = pml.Dataset.from_device_pointers(ctx, n, D_FE, fe_ptr.data_ptr(), fe_cols.data_ptr(), ones.data_ptr(),
                                                 pml.abi.F32, label_dev.data_ptr())
        ctx.synchronize()
        t_pack_fe = time.perf_counter() - t0
        del fe_cols, fe_ptr, ones, cdf
        torch.cuda.empty_cache()
        peer = pml.Peer(ctx, D_FE) if world > 1 else None
        passes = 2

        def all_sum(x):
            t = torch.tensor([x], dtype=torch.float64, device=device)
            dist.all_reduce(t)
            return float(t)

        coords = {"fixed": FixedEffect(fe_ds, pml.LOGISTIC, l2RegWeight=1.0, peer=peer),
                  "per-member": RandomEffect(batch, None, pml.LOGISTIC, l2RegWeight=1.0)}
        descent = CoordinateDescent(coords, n, descent_iterations=passes, all_reduce_sum=all_sum if world > 1 else None)
        warm_up_coordinates(torch, dist, world, coords, 1)
        barrier()
        t0 = time.perf_counter()
        _, history = descent.run()
        barrier()
        cd_wall = time.perf_counter() - t0
        cd = {"wall_s": cd_wall, "passes": passes, "fixed_effect": {"d": D_FE, "entries_per_row": K_FE, "pack_s_rank0": round(t_pack_fe, 2),
                                                                    "device_GB_rank0": fe_ds.info().device_bytes / 1e9},
              "updates_rank0": [{k: (round(v, 6) if isinstance(v, float) else v) for k, v in h.items()} for h in history]}
        if peer is not None:
            peer.close()
        fe_ds.close()
    stats = torch.tensor([eval_ms, wall, float(n), float(len(mine)), float(sinfo["iterations"].sum()),
                          cd["wall_s"] if cd else 0.0], dtype=torch.float64, device=device)
    mx = stats.clone()
    if world > 1:
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
    info = batch.dataset.info()
    if rank == 0:
        peak, peak_src = peaks()
        eval_ms, wall = float(mx[0]), float(mx[1])
        rows_total, ents_total = int(stats[2]), int(stats[3])
        nbytes = batch_bytes(len(mine), n, n * 10, batch.d_total, info.value_bytes)
        emit({
            "metric": "per-entity problems solved/s (lock-step TRON, L2 logistic)", "value": ents_total / wall,
            "unit": "entities/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": eval_ms,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "configs[4] random-effect coordinate: %d per-member problems (%d rows, 10 entries per row from a "
                                   "200-dim member-side shard, index-map projected per member), entities sharded over ranks, "
                                   "no collective" % (ents_total, rows_total),
                       "entities_rank0": int(len(mine)), "rows_rank0": n, "rows_max_over_ranks": int(mx[2]),
                       "partitioner": "bin packing of the 10000 heaviest entities + id mod N (RandomEffectDataSetPartitioner.scala:113-146)",
                       "pack_s_rank0": round(t_pack, 2)},
            "solve": {"wall_s_max_over_ranks": wall, "mean_iterations_per_entity": float(stats[4]) / ents_total,
                      "batched_value_gradient_evals_rank0": sinfo["evaluations"],
                      "batched_hessian_vector_evals_rank0": sinfo["hessian_vector_evaluations"],
                      "reasons_rank0": {REASONS[c]: int((sinfo["reasons"] == c).sum()) for c in range(len(REASONS))}},
            "batched_evaluation": {"ms_max_over_ranks": eval_ms, "entity_evaluations_per_s": ents_total / (eval_ms * 1e-3)},
            "roofline": {"bound": "hbm", "kernel": "pml_batch_value_and_gradient_dev (rank 0's shard)",
                         "algorithmic_bytes": int(nbytes), "achieved": nbytes / (eval_ms * 1e-3) / 1e9, "peak": peak,
                         "unit": "GB/s", "frac": nbytes / (eval_ms * 1e-3) / 1e9 / peak, "peak_source": peak_src, "traffic": None},
            "coordinate_descent": None if cd is None else dict(
                cd, wall_s=float(mx[5]), rows_total=rows_total, fixed_effect_entries_total=rows_total * 25,
                row_passes_per_s=rows_total * cd["passes"] / float(mx[5])),
            "gpu_launches": int(ctx.launch_count),
        })
    batch.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


C2_FEATS = 10_000_000
C2_ROWS_PER_GPU = 25_000_000
C2_NNZ_PER_ROW = 39
C2_GEN_CHUNK = 500_000
# Above this many coefficients an 8-rank exchange goes through NCCL's in-switch all-reduce (+ finalize_dev) instead of
# the peer-memory merge kernel: measured (profiles/r01_merge_latency.txt) 322 vs 275 us at d = 8 M on 8 GPUs, while the
# peer kernel wins at every size on 2 and 4 GPUs and below ~4 M coefficients on 8.
PEER_MERGE_MAX_D_AT_8 = 4_000_000


def run_config2(args):
    """configs[2]: Criteo-shaped sparse logistic regression, 200 M rows x 10 M hashed features on 8 GPUs -- 25 M rows
    per GPU whatever N is (weak scaling): 39 entries per row (13 from a small dense range + 26 hashed Zipf(1.05)
    categoricals), values 1.0 (stored as indices only), CTR ~3 %.  A step = one value+gradient evaluation of all
    ranks' rows with one exchange of d+2 doubles; also the OWL-QN solve the configuration names (alpha = 1 and 0.5)."""
    import torch
    import torch.distributed as dist
    import photon_ml_b200 as pml
    from photon_ml_b200.function import DeviceGLMLossFunction, NoNormalization
    from photon_ml_b200.optimization import OWLQN
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    n_local = int(os.environ.get("PML_BENCH_ROWS_PER_GPU", C2_ROWS_PER_GPU))
    D = C2_FEATS
    ctx = pml.Context(local_rank)
    ext = torch.cuda.ExternalStream(ctx.stream, device=device)
    p = np.arange(1, D - 64 + 1, dtype=np.float64) ** -1.05
    cdf = torch.from_numpy(np.cumsum(p / p.sum())).to(device)      # built on the host: the same bits on every rank
    g = torch.Generator(device=device)
    g.manual_seed(3_000 + rank)
    t0 = time.time()
    cols, labels = [], []
    for _ in range(n_local // C2_GEN_CHUNK):
        dense = torch.randint(0, 64, (C2_GEN_CHUNK, 13), generator=g, device=device)
        hashed = 64 + torch.searchsorted(cdf, torch.rand((C2_GEN_CHUNK, 26), generator=g, device=device,
                                                         dtype=torch.float64)).clamp_(max=D - 65)
        cc, _ = torch.sort(torch.cat([dense, hashed], 1), dim=1)
        cols.append(cc.to(torch.int32).reshape(-1))
        labels.append((torch.rand(C2_GEN_CHUNK, generator=g, device=device) < 0.03).to(torch.float64))
    cols, labels = torch.cat(cols), torch.cat(labels)
    n_local = labels.numel()
    vals = torch.ones(cols.numel(), dtype=torch.float32, device=device)
    row_ptr = torch.arange(n_local + 1, dtype=torch.int64, device=device) * C2_NNZ_PER_ROW
    torch.cuda.synchronize()
    t_gen = time.time() - t0
    t0 = time.time()
    ds = pml.Dataset.from_device_pointers(ctx, n_local, D, row_ptr.data_ptr(), cols.data_ptr(), vals.data_ptr(), pml.abi.F32,
                                          labels.data_ptr(), storage=pml.STORE_AUTO)
    ctx.synchronize()
    t_pack = time.time() - t0
    del cols, vals, row_ptr, labels, cdf
    torch.cuda.empty_cache()
    info = ds.info()
    coef = torch.randn(D, generator=g, device=device, dtype=torch.float64) * 0.05
    if world > 1:
        dist.broadcast(coef, 0)
    raw = torch.empty(D + 2, dtype=torch.float64, device=device)
    out_value = torch.empty(1, dtype=torch.float64, device=device)
    out_grad = torch.empty(D, dtype=torch.float64, device=device)
    peer, host_peer, collective = None, None, "none"
    if world > 1:
        want_peer = os.environ.get("PML_BENCH_COLLECTIVE", "auto")
        use_peer = want_peer == "peer" or (want_peer == "auto" and not (world >= 8 and D > PEER_MERGE_MAX_D_AT_8))
        collective = "nccl all-reduce of d+2 doubles, then finalize_dev (chosen by size: d = %d on %d ranks)" % (D, world)
        # the peer handle exists either way: the end-to-end step below goes through the host-buffer exchange
        # (pml_value_and_gradient_allreduce), whose slice-wise PCIe traffic matters far more at d = 10 M than the
        # 50 us between the two merges
        host_peer = pml.Peer(ctx, D)
        if use_peer:
            peer = host_peer
            collective = "peer_merge_kernel: sums over NVLink peer memory in rank order, fused with getVector"
            result = torch.as_tensor(peer.result_array(), device=device)
            out_grad, out_value = result[:D], result[D:D + 1]

    def step(w=coef):
        if peer is not None:
            peer.value_and_gradient(ds, pml.LOGISTIC, w.data_ptr())
            return
        ds.value_and_gradient_raw_dev(pml.LOGISTIC, w.data_ptr(), raw.data_ptr())
        if world > 1:
            dist.all_reduce(raw)
        ds.finalize_dev(pml.LOGISTIC, raw.data_ptr(), out_value.data_ptr(), out_grad.data_ptr())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    class ShardedObjective(DeviceGLMLossFunction):
        """calculate() over all ranks' rows: every rank runs the same optimizer on the same (bit-identical) results."""

        def calculate(self, data, coefficients, normalizationContext, want_gradient=True):
            step(coefficients)
            f = float(out_value)
            return f, (out_grad.clone() if want_gradient else None)

    with torch.cuda.stream(ext):
        for _ in range(args.warmup):
            step()
        barrier()
        sampler = ClockSampler(local_rank)
        if rank == 0:
            sampler.start()
        launches0 = ctx.launch_count
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        a.record(ext)
        for _ in range(args.steps):
            step()
        b.record(ext)
        barrier()
        ms = a.elapsed_time(b)
        launches = ctx.launch_count - launches0
        clocks = sampler.stop() if rank == 0 else None
        ctx.read_profile()
        ctx.set_profiling(True)
        for _ in range(3):
            step()
        barrier()
        profile = ctx.read_profile()
        ctx.set_profiling(False)
        # end to end: host coefficients in, host gradient out (every rank its slice when the peer exchange is in use)
        h_coef = torch.empty(D, dtype=torch.float64).pin_memory()
        h_coef.copy_(coef.cpu())
        h_grad = torch.empty(D, dtype=torch.float64).pin_memory()
        h_coef_np, h_grad_np = h_coef.numpy(), h_grad.numpy()
        shared, e2e_iter = None, [0]
        if world > 1:      # one host buffer for all ranks, as in the configs[1] run: the driver ends up with the full gradient
            shared = SharedHostVectors(torch, dist, rank, world, D, coef.cpu().numpy())
            h_coef_np, h_grad_np = shared.coef, shared.grad

        def e2e_step():
            if world == 1:
                return ds.value_and_gradient(pml.LOGISTIC, h_coef_np, out=h_grad_np)[0]
            value = host_peer.value_and_gradient_host(ds, pml.LOGISTIC, h_coef_np, h_grad_np)
            e2e_iter[0] += 1
            shared.arrive_and_wait(e2e_iter[0])
            return value

        for _ in range(3):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_value = e2e_step()
        barrier()
        e2e_s = time.perf_counter() - t0
        if shared is not None:
            shared.close()
        # the configuration's optimizer: OWL-QN, L1 (alpha = 1) and elastic net (alpha = 0.5; the reference mixes the L2
        # part in only for RegularizationType.L2, so ELASTIC_NET gets the L1 part alone: DistributedGLMLossFunction.scala:160-169)
        owlqn = {}
        if not args.no_convergence:
            for alpha in (1.0, 0.5):
                opt = OWLQN(alpha * 1.0, NoNormalization(), maxNumIterations=int(os.environ.get("PML_BENCH_OWLQN_ITERS", "30")))
                obj = ShardedObjective(ctx, pml.LOGISTIC)
                barrier()
                t0 = time.perf_counter()
                w, f = opt.optimize(obj, torch.zeros(D, dtype=torch.float64, device=device), ds)
                barrier()
                owlqn["alpha_%g" % alpha] = {"wall_s": time.perf_counter() - t0, "iterations": opt.currentState.iter,
                                             "evaluations": opt.evaluations, "reason": opt.convergenceReason, "objective": f,
                                             "nonzero_coefficients": int((w != 0).sum())}
    times = torch.tensor([ms, e2e_s], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms, e2e_s = float(times[0]), float(times[1])
    if rank == 0:
        peak, peak_src = peaks()
        per = lambda k: profile[k][0] / max(1, profile[k][1])   # noqa: E731
        ms_per_step = ms / args.steps
        V = info.value_bytes
        n, nnz = info.n, info.nnz
        step_bytes = 2 * nnz * (V + 4) + 28 * n + 28 * D       # per GPU: two entry streams, label + r written and read, w, g
        row_bytes = nnz * (V + 4) + (n + 1) * 4 + n * 16 + D * 8
        col_bytes = nnz * (V + 4) + n * 8 + D * 12
        row_ms = per("row_pass") + per("row_epilogue")
        col_ms = per("col_pass") + per("col_pass_tiled")
        dom_bytes, dom_ms, dom = ((col_bytes, col_ms, "CSC pass X^T r (tiled hot columns + cold columns in super blocks)")
                                  if col_ms >= row_ms else (row_bytes, row_ms, "CSR pass X w (seg_spmv_warp_kernel + row epilogue)"))
        emit({
            "metric": "value+gradient evals/s", "value": 1e3 / ms_per_step, "unit": "evals/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "configs[2]: Criteo-shaped logistic, %d rows per GPU x %d hashed features, %d entries per row, "
                                   "indicator values stored as indices only" % (n, D, C2_NNZ_PER_ROW),
                       "rows_total": int(n) * world, "rows_per_gpu": int(n), "nnz_per_gpu": int(nnz), "d": D,
                       "value_bytes_per_entry": int(V), "collective": collective,
                       "l2_flush": "inputs larger than L2 (%.1f GB streamed per step per GPU)" % (step_bytes / 1e9),
                       "csc_tiling": {"hot_cols": int(info.tiled_hot_cols), "tiles": int(info.tiled_tiles),
                                      "tile_rows": int(info.tiled_tile_rows)},
                       "device_GB": info.device_bytes / 1e9, "generate_s": round(t_gen, 1), "pack_s": round(t_pack, 2)},
            "rows_per_s": int(n) * world / (ms_per_step * 1e-3),
            "e2e": {"value": args.steps / e2e_s, "unit": "evals/s", "h2d_bytes_per_step": D * 8,
                    "d2h_bytes_per_step": D * 8 + 8 * world, "ms_per_step": 1e3 * e2e_s / args.steps, "objective": e2e_value,
                    "api": "pml_value_and_gradient (host buffers, pinned)" if world == 1 else
                           "pml_value_and_gradient_allreduce: every rank uploads 1/N of the coefficients and downloads 1/N of "
                           "the gradient over its own PCIe link (one host copy of the vector per rank: Spark's broadcast)"},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": dom_bytes / (dom_ms * 1e-3) / 1e9, "peak": peak,
                         "unit": "GB/s", "frac": dom_bytes / (dom_ms * 1e-3) / 1e9 / peak, "peak_source": peak_src,
                         "traffic": None, "algorithmic_bytes_per_launch": int(dom_bytes), "avg_launch_ms": dom_ms},
            "step": {"algorithmic_bytes_per_gpu": int(step_bytes), "GBps_per_gpu": step_bytes / (ms_per_step * 1e-3) / 1e9,
                     "frac": step_bytes / (ms_per_step * 1e-3) / 1e9 / peak},
            "kernels": {k: {"ms_per_launch": v[0] / max(1, v[1]), "launches": int(v[1])} for k, v in profile.items()},
            "owlqn_logistic_lambda_1": owlqn,
        })
    ds.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def warm_up_coordinates(torch, dist, world, coords, n_scores):
    """Untimed: the first evaluations of every kind allocate lazily, capture their CUDA graphs and (N > 1) set up the
    collectives -- none of which belongs to a coordinate-descent pass.  The models are not touched."""
    from photon_ml_b200.batched import BatchedTRON
    from photon_ml_b200.function import NoNormalization
    for c in coords.values():
        if hasattr(c, "batch"):
            BatchedTRON(c.batch, c.loss, l2RegWeight=c.l2, maxNumIterations=1).optimize()
        else:
            with c.objective.stream():
                zero = torch.zeros_like(c.model)
                c.objective.calculate(c.ds, zero, NoNormalization())
                c.objective.hessianVector(c.ds, zero, zero, NoNormalization())
                c.score_local()
    if world > 1:
        dist.all_reduce(torch.zeros(n_scores, dtype=torch.float64, device=next(iter(coords.values())).device))
    torch.cuda.synchronize()


def run_config3(args):
    """configs[3]: the GLMix MovieLens-20M shape -- fixed effect (d = 1000, 30 indicator features per row) + per-user
    random effect (138 493 users, 5 of 20 movie-side indicators + intercept) + per-movie random effect (27 278 movies,
    6 of 30 user-side indicators + intercept), logistic loss, L2 = 1, 3 coordinate-descent passes fixed -> per-user ->
    per-movie.  One process per GPU: users are bin-packed over the ranks and a rank owns the rows of ITS users for the
    fixed effect (merged through the peer exchange) and the per-user effect; movies are bin-packed separately, so the
    per-movie coordinate's rows live elsewhere and the global score vectors are summed over the ranks after every
    update (the reference joins score RDDs by row id).  value = rows x passes per second, whole job."""
    import torch
    import torch.distributed as dist
    import photon_ml_b200 as pml
    from photon_ml_b200 import parallel
    from photon_ml_b200.game import CoordinateDescent, FixedEffect, RandomEffect
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    scale = float(os.environ.get("PML_BENCH_SCALE", "1.0"))
    n_users, n_movies, d_fe = int(138_493 * scale), int(27_278 * scale), 1000
    ctx = pml.Context(local_rank)
    gen = torch.Generator(device=dev)
    gen.manual_seed(4)                       # every rank draws the same 20 M rows, then keeps its own

    def indicators(n, k, d, intercept):
        cols, _ = torch.sort(torch.randint(0, d, (n, k), generator=gen, device=dev, dtype=torch.int32), dim=1)
        if intercept:
            cols = torch.cat([cols, torch.full((n, 1), d, device=dev, dtype=torch.int32)], dim=1)
        return cols

    t0 = time.time()
    rows = torch.exp(torch.randn(n_users, generator=gen, device=dev, dtype=torch.float64) + np.log(144.0) - 0.5)
    rows = rows.clamp_(min=20).to(torch.int64)
    n = int(rows.sum())
    user_row_ptr = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), torch.cumsum(rows, 0)])
    pz = np.arange(1, n_movies + 1, dtype=np.float64) ** -0.9
    cdf = torch.from_numpy(np.cumsum(pz / pz.sum())).to(dev)
    movie_of_row = torch.searchsorted(cdf, torch.rand(n, generator=gen, device=dev, dtype=torch.float64)).clamp_(max=n_movies - 1)
    movie_map = torch.argsort(movie_of_row, stable=True)                 # per-movie row i = global row movie_map[i]
    movie_rows = torch.bincount(movie_of_row, minlength=n_movies)
    movie_row_ptr = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), torch.cumsum(movie_rows, 0)])
    label = (torch.rand(n, generator=gen, device=dev) < 0.5).to(torch.float64)
    fe_cols = indicators(n, 30, d_fe, False)
    user_cols = indicators(n, 5, 20, True)                               # rows are grouped by user
    movie_cols = indicators(n, 6, 30, True)                              # generated in per-movie order
    torch.cuda.synchronize()
    t_gen = time.time() - t0

    def owned(ent_rows, ent_ptr):
        """Entities of this rank (the reference's partitioner rule) and the positions of their rows, entity-major."""
        assign = torch.from_numpy(parallel.shard_entities(ent_rows.cpu().numpy(), world, capacity=10000)).to(dev)
        mine = torch.nonzero(assign == rank).flatten()
        lens = ent_rows[mine]
        start = ent_ptr[mine]
        local_ptr = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), torch.cumsum(lens, 0)])
        pos = torch.repeat_interleave(start - local_ptr[:-1], lens) + torch.arange(int(lens.sum()), device=dev)
        return mine, local_ptr, pos

    t0 = time.time()
    my_users, my_user_ptr, my_rows = owned(rows, user_row_ptr)           # positions = global row ids (user-major order)
    my_movies, my_movie_ptr, my_pos = owned(movie_rows, movie_row_ptr)   # positions in the per-movie order
    movie_global = movie_map[my_pos]
    n_fe = my_rows.numel()
    fe_local = fe_cols[my_rows].contiguous()
    fe_ptr = torch.arange(n_fe + 1, dtype=torch.int64, device=dev) * 30
    ones = torch.ones(n_fe * 30, dtype=torch.float32, device=dev)
    fe_label = label[my_rows].contiguous()
    fe_ds = pml.Dataset.from_device_pointers(ctx, n_fe, d_fe, fe_ptr.data_ptr(), fe_local.data_ptr(), ones.data_ptr(),
                                             pml.abi.F32, fe_label.data_ptr())
    ctx.synchronize()
    user_batch = pml.Batch(ctx, my_user_ptr.cpu().numpy(), np.full(my_users.numel(), 21, np.int32),
                           np.arange(n_fe + 1, dtype=np.int64) * 6, user_cols[my_rows].cpu().numpy().ravel(),
                           np.ones(n_fe * 6, np.float32), fe_label.cpu().numpy())
    n_mv = my_pos.numel()
    movie_batch = pml.Batch(ctx, my_movie_ptr.cpu().numpy(), np.full(my_movies.numel(), 31, np.int32),
                            np.arange(n_mv + 1, dtype=np.int64) * 7, movie_cols[my_pos].cpu().numpy().ravel(),
                            np.ones(n_mv * 7, np.float32), label[movie_global].cpu().numpy())
    ctx.synchronize()
    t_pack = time.time() - t0
    del fe_cols, user_cols, movie_cols, fe_local, ones, fe_ptr, movie_of_row, movie_map, cdf
    torch.cuda.empty_cache()
    peer = pml.Peer(ctx, d_fe) if world > 1 else None

    def all_sum(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        return float(t)

    passes = 3
    coords = {"fixed": FixedEffect(fe_ds, pml.LOGISTIC, l2RegWeight=1.0, row_map=my_rows, peer=peer),
              "per-user": RandomEffect(user_batch, my_rows, pml.LOGISTIC, l2RegWeight=1.0),
              "per-movie": RandomEffect(movie_batch, movie_global, pml.LOGISTIC, l2RegWeight=1.0)}
    cd = CoordinateDescent(coords, n, descent_iterations=passes, all_reduce_sum=all_sum if world > 1 else None,
                           exchange_scores=(lambda t: dist.all_reduce(t)) if world > 1 else None)
    warm_up_coordinates(torch, dist, world, coords, n)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    _, history = cd.run()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    wall = time.perf_counter() - t0
    stats = torch.tensor([wall, float(n_fe), float(n_mv)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
    if rank == 0:
        wall = float(stats[0])
        decreasing = all(history[i + 1]["objective"] <= history[i]["objective"] * (1 + 1e-12) for i in range(len(history) - 1))
        emit({
            "metric": "coordinate-descent throughput (rows x passes per second)", "value": n * passes / wall,
            "unit": "row-passes/s", "n_gpus": world, "steps": passes, "warmup": 0, "ms_per_step": 1e3 * wall / passes,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "configs[3]: GLMix MovieLens-20M shape, %d rows, %d users, %d movies; fixed effect d = %d + per-user "
                                   "+ per-movie random effects, logistic, L2 = 1, %d coordinate-descent passes" %
                                   (n, n_users, n_movies, d_fe, passes),
                       "rows_total": n, "fixed_and_user_rows_max_over_ranks": int(stats[1]), "movie_rows_max_over_ranks": int(stats[2]),
                       "exchange": "fixed effect: peer-memory merge of d+2 sums per evaluation; scores: one all-reduce of the "
                                   "global score vector (%d doubles) per coordinate update" % n if world > 1 else "none",
                       "generate_s": round(t_gen, 1), "pack_s_rank0": round(t_pack, 1)},
            "wall_s": wall, "objective_decreasing": decreasing,
            "updates": [{k: (round(v, 6) if isinstance(v, float) else v) for k, v in h.items()} for h in history],
            "gpu_launches": int(ctx.launch_count),
        })
    if peer is not None:
        peer.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def full_rows_on_host():
    """All N_ROWS rows of the bench workload as host arrays.  Generated on the GPU when there is one (the very same
    generator and seeds as the GPU arm, so both arms evaluate identical rows) and copied down; on a box without a GPU
    the same distribution is drawn with torch's CPU generator."""
    import torch
    device = torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")
    rows = generate_rows(torch, device, range(N_ROWS // GEN_CHUNK))
    checksum = data_checksum(torch, *rows)
    host = rows_to_host(rows)
    del rows
    if device.type == "cuda":
        torch.cuda.empty_cache()
    return host, checksum, device.type


def cpu_baseline(host_rows, steps, warmup):
    """The oracle (CPU restatement of the reference's row loop, partitioned the way Spark local[N] runs
    ValueAndGradientAggregator.calculateValueAndGradient, :235-250: one partition per core, a dense d-vector per
    partition, pairwise merge) timed on this box's host cores over ALL rows of the workload -- measured, not scaled."""
    import oracle
    cores = oracle.max_threads()
    row_ptr, cols, vals, labels = host_rows
    n = len(labels)
    data = oracle.Data(n, N_FEATS, row_ptr, cols, vals, labels)
    coef = np.random.default_rng(99).normal(scale=0.02, size=N_FEATS)
    for _ in range(warmup):
        oracle.value_and_gradient(data, oracle.POISSON, coef, parts=cores)
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        oracle.value_and_gradient(data, oracle.POISSON, coef, parts=cores)
        times.append(time.perf_counter() - t0)
    per_eval = statistics.median(times)
    return {"value": 1.0 / per_eval, "unit": "evals/s", "cores": cores, "kind": "port",
            "sample": "full: all %d rows (%d nnz), %d partitions/threads, %d warm-up + %d timed evaluations, median %.4f s "
                      "(mean %.4f s). CPU restatement of the reference loop (not Spark: no JVM on this box)"
                      % (n, len(cols), cores, warmup, steps, per_eval, statistics.mean(times)),
            "eval_s_median": per_eval, "eval_s_mean": statistics.mean(times), "eval_s_total": sum(times)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, args.steps)
    warmup = max(1, args.warmup)
    host_rows, checksum, where = full_rows_on_host()
    base = cpu_baseline(host_rows, steps=steps, warmup=warmup)
    ms_per_step = 1e3 * base["eval_s_total"] / steps    # the timed region's own clock: K evaluations back to back
    value = 1e3 / ms_per_step
    base["value"] = value
    out = {
        "impl": "reference",
        "metric": "value+gradient evals/s", "value": value, "unit": "evals/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "configs[1]: Poisson regression, 10M rows x 1M feats, 66 Zipf(1.1) draws/row "
                               "(%.1f stored nnz/row), fp32 values, fp64 accumulation" % (checksum[1] / checksum[0]),
                   "reference_arm": "the same rows (values widened to the Doubles the reference holds); each step = one "
                                    "evaluation of ALL rows on the host cores, %d partitions / threads" % base["cores"],
                   "rows_total": N_ROWS, "d": N_FEATS, "rows_generated_on": where,
                   "data_checksum": dict(zip(("rows", "nnz", "sum_col", "sum_val", "sum_label"), checksum))},
        "cpu_baseline": base,
        "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(out)


_REAL_STDOUT = None


def emit(obj):
    """The one JSON line, on the process's real stdout."""
    line = (json.dumps(obj) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, line)


def main():
    # Native libraries (NCCL's version banner, ...) write to fd 1: park everything but the JSON line on stderr.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-convergence", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--config", type=int, default=1, choices=[1, 2, 3, 4],
                    help="BASELINE.json configs index: 1 = the headline workload (default); 2 = Criteo-shaped logistic, "
                         "3 = GLMix coordinate descent (fixed + per-user + per-movie effects), "
                         "25 M rows per GPU (weak scaling); 4 = the random-effect coordinate, entities sharded over GPUs")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    elif args.config == 4:
        run_config4(args)
    elif args.config == 2:
        run_config2(args)
    elif args.config == 3:
        run_config3(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
