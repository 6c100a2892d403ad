#!/usr/bin/env python
"""2+ GPU probe: throughput of multimem.red (in-switch reduction) vs local red, and of the wgrad kernel
with a multicast destination.  torchrun this."""
import ctypes
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fdt.ops import _lib  # noqa: E402
from fdt.ops._lib import call, i64  # noqa: E402


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e3


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from fdt.parallel.symm import SymmetricHeap
    heap = SymmetricHeap.create(dev)
    n = 16 * 1024 * 1024
    buf = heap.alloc_f32(n)
    mc = heap.multicast_addr(buf)
    src = torch.randn(n, device=dev)
    t_mc = timed(lambda: call("fdt_comm_mc_red_add", ctypes.c_void_p(mc), src, i64(n), 1.0))
    if rank == 0:
        print(f"multimem.red of {n*4/1e6:.0f} MB: {t_mc:.1f} us -> {n*4/t_mc/1e3:.1f} GB/s per GPU", flush=True)
    # wgrad: local vs multicast destination
    for (N, C, O, HW, k) in [(128, 64, 64, 32, 3), (128, 512, 512, 4, 3), (128, 256, 1024, 8, 1)]:
        x = torch.randn(N, C, HW, HW, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
        dy = torch.randn(N, O, HW, HW, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
        loc = torch.zeros(O * k * k * C, device=dev)
        t_loc = timed(lambda: call("fdt_conv_wgrad", dy, x, loc, N, HW, HW, C, O, k, k // 2, 1.0, 0))
        t_mcw = timed(lambda: call("fdt_conv_wgrad", dy, x, ctypes.c_void_p(mc), N, HW, HW, C, O, k, k // 2, 0.5, 1))
        if rank == 0:
            print(f"wgrad C{C} O{O} {HW}x{HW} k{k}: local {t_loc:.1f} us | multicast {t_mcw:.1f} us", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
