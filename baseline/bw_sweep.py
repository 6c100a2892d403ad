#!/usr/bin/env python
"""Gradient all-reduce / reduce-scatter / all-gather sweep, 1 KB – 1 GB (BASELINE.json config 5):
this framework's fused NVLink kernels (P2P and NVLS-multicast paths) against NCCL through
torch.distributed, on the same buffers.  Launch with torchrun; device-timed (CUDA events), max over
ranks, 10 timed iterations after 3 warm-ups.

"busbw" uses the NCCL convention: all-reduce 2(n-1)/n · bytes / t; reduce-scatter / all-gather
(n-1)/n · bytes / t, bytes = full buffer size.  Note the fused kernels do extra work in the same
pass (1/world scaling, Σg² for clipping; bf16 shadow cast for all-gather)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(fn, world, iters=10, warm=3):
    for _ in range(warm):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    evs = []
    for _ in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record()
        evs.append((s, e))
    torch.cuda.synchronize()
    ms = sorted(s.elapsed_time(e) for s, e in evs)[len(evs) // 2]
    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from fdt.parallel.symm import SymmetricHeap
    heap = SymmetricHeap.create(dev)
    sizes = [1 << k for k in range(10, 31, 2)] + [94 * 1024 * 1024, 117 * 1024 * 1024]   # + DDP / FSDP payloads
    max_bytes = int(os.environ.get("SWEEP_MAX_BYTES", str(1 << 30)))
    rows = []
    for nbytes in sorted(s for s in sizes if s <= max_bytes):
        n = max(nbytes // 4, 4 * world)
        n = n // (4 * world) * (4 * world)
        buf = heap.alloc_f32(n)
        ref = torch.zeros(n, device=dev)
        shadow = torch.empty(n, dtype=torch.bfloat16, device=dev)
        region = n // world
        lo, hi = rank * region, (rank + 1) * region
        has_mc = bool(heap._ctx(buf).mc)
        row = {"bytes": n * 4}
        for mc in ([True, False] if has_mc else [False]):
            heap.use_multicast = mc
            tag = "nvls" if mc else "p2p"
            row[f"ar_{tag}_us"] = timed(lambda: heap.all_reduce_mean_(buf), world) * 1e3
            row[f"rs_{tag}_us"] = timed(lambda: heap.reduce_scatter_mean_(buf, lo, hi), world) * 1e3
        row["ag_fused_us"] = timed(lambda: heap.all_gather_(buf, lo, hi, shadow=shadow), world) * 1e3
        row["ar_nccl_us"] = timed(lambda: dist.all_reduce(ref), world) * 1e3
        row["rs_nccl_us"] = timed(lambda: dist.reduce_scatter_tensor(ref[lo:hi], ref), world) * 1e3
        row["ag_nccl_us"] = timed(lambda: dist.all_gather_into_tensor(ref, ref[lo:hi]), world) * 1e3
        rows.append(row)
        del buf
        if rank == 0:
            b = row["bytes"]
            f = 2 * (world - 1) / world
            best = min(row.get("ar_nvls_us", 1e30), row["ar_p2p_us"])
            print(f"{b:>11d} B | all-reduce fused {best:9.1f} us ({f * b / best / 1e3:7.1f} GB/s busbw) "
                  f"nccl {row['ar_nccl_us']:9.1f} us ({f * b / row['ar_nccl_us'] / 1e3:7.1f}) | "
                  f"rs fused {min(row.get('rs_nvls_us', 1e30), row['rs_p2p_us']):8.1f} nccl {row['rs_nccl_us']:8.1f} | "
                  f"ag fused(+cast) {row['ag_fused_us']:8.1f} nccl {row['ag_nccl_us']:8.1f}", flush=True)
    if rank == 0:
        print(json.dumps({"world": world, "rows": rows}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
