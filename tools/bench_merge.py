"""Latency of the merge step alone (not part of bench.py): the library's peer-memory kernel (NVSwitch multicast or
peer-to-peer) against torch's NCCL all-reduce of the d+2 raw sums followed by finalize_dev.  Launch with torchrun."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import photon_ml_b200 as pml  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ctx = pml.Context(local)
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    out = {}
    for d in (1 << 12, 1 << 16, 1 << 20, 1 << 23):
        rng = np.random.default_rng(rank)
        n = 64
        row_ptr = np.arange(n + 1, dtype=np.int64) * 4
        col = rng.integers(0, d, size=4 * n).astype(np.int32)
        col.reshape(n, 4).sort(axis=1)
        ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, np.ones(4 * n, np.float32), np.zeros(n))
        raw = torch.randn(d + 2, dtype=torch.float64, device=dev)
        res = torch.empty(d, dtype=torch.float64, device=dev)
        val = torch.empty(1, dtype=torch.float64, device=dev)
        row = {}
        for mode in ("multicast", "p2p", "nccl"):
            peer = None
            if mode != "nccl":
                peer = pml.Peer(ctx, d, multicast=(mode == "multicast"))
                if peer.transport != mode:
                    peer.close()
                    continue
                torch.as_tensor(pml.abi._DeviceArray(peer.raw_ptr, d + 2, peer), device=dev).copy_(raw)

            def step():
                if peer is not None:
                    peer.merge(ds, pml.SQUARED)
                else:
                    dist.all_reduce(raw)
                    ds.finalize_dev(pml.SQUARED, raw.data_ptr(), val.data_ptr(), res.data_ptr())
            with torch.cuda.stream(ext):
                for _ in range(20):
                    step()
                torch.cuda.synchronize()
                dist.barrier()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                iters = 200
                a.record(ext)
                for _ in range(iters):
                    step()
                b.record(ext)
                torch.cuda.synchronize()
                t = torch.tensor([a.elapsed_time(b) / iters * 1e3], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                row[mode] = round(float(t), 2)
            if peer is not None:
                peer.close()
            raw.normal_()
        ds.close()
        out[d] = row
    if rank == 0:
        print("merge latency in us per call, max over %d ranks: %s" % (world, out))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
