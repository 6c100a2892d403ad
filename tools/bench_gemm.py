#!/usr/bin/env python
"""Micro-benchmark: tcgen05 GEMM / conv kernels vs cuBLAS / cuDNN, and the channel-wise kernels, on
the ResNet-50 and transformer shapes.  Kernel-only time: each measurement captures NB back-to-back
calls on NB different buffer sets (so inputs are not L2-resident across calls: NB x footprint > L2)
in a CUDA graph and times graph replays with CUDA events."""
import math
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fdt.ops import gemm_sm100 as g  # noqa: E402
from fdt.ops._lib import call, i64  # noqa: E402


def graph_time(fns, reps=10):
    """fns: list of callables (one per buffer set). Returns median us per call."""
    for f in fns:
        f()
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for f in fns:
            f()
    gr.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); gr.replay(); e.record()
        e.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2] * 1e3 / len(fns)


def gemm_table():
    print("gemm: us 1-CTA kernel | us 2-CTA kernel (cta_group::2) | us cublas | TFLOP/s 1-CTA | TFLOP/s 2-CTA | TFLOP/s cublas")
    for (M, N, K) in [(8192, 512, 512), (8192, 1536, 512), (8192, 1024, 512), (8192, 512, 1024), (131072, 256, 64),
                      (4096, 4096, 4096), (8192, 8192, 8192)]:
        nb = 2 if M * N > 5e7 else 6
        As = [torch.randn(M, K, device="cuda").bfloat16() for _ in range(nb)]
        b = torch.randn(N, K, device="cuda").bfloat16()
        outs = [torch.empty(M, N, dtype=torch.bfloat16, device="cuda") for _ in As]
        tn = graph_time([(lambda a=a, o=o: call("fdt_gemm_nt_ex", a, b, o, None, None, None, None, M, N, K, 0)) for a, o in zip(As, outs)])
        t2 = graph_time([(lambda a=a: g.gemm2_nt(a, b)) for a in As])
        tl = graph_time([(lambda a=a: F.linear(a, b)) for a in As])
        fl = 2 * M * N * K / 1e6
        print(f"M{M} N{N} K{K}: {tn:8.1f} | {t2:8.1f} | {tl:8.1f} | {fl / tn:7.1f} | {fl / t2:7.1f} | {fl / tl:7.1f}")


def main():
    torch.backends.cudnn.benchmark = True
    if "--gemm-only" in sys.argv:
        return gemm_table()
    B = 128
    print("conv fwd (N=128): us native | us native + BN statistics epilogue | us cudnn | TFLOP/s native | GB/s native (min traffic)")
    for (C, O, HW, k) in [(64, 64, 32, 3), (64, 64, 32, 1), (64, 256, 32, 1), (256, 64, 32, 1), (128, 128, 16, 3), (128, 512, 16, 1),
                          (512, 128, 16, 1), (256, 256, 8, 3), (256, 1024, 8, 1), (1024, 256, 8, 1), (512, 512, 4, 3),
                          (512, 2048, 4, 1), (2048, 512, 4, 1)]:
        by = (B * C * HW * HW + O * C * k * k + B * O * HW * HW) * 2
        nb = max(2, min(8, int(160e6 // by) + 1))
        xs = [torch.randn(B, C, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last) for _ in range(nb)]
        w = (torch.randn(O, C, k, k, device="cuda") / math.sqrt(C * k * k)).bfloat16().contiguous(memory_format=torch.channels_last)
        tn = graph_time([(lambda x=x: g.conv_fwd(x, w, 1, k // 2)) for x in xs])
        sums = torch.zeros(2 * O, device="cuda")
        ts = graph_time([(lambda x=x: g.conv_fwd(x, w, 1, k // 2, sums)) for x in xs])
        tl = graph_time([(lambda x=x: F.conv2d(x, w, padding=k // 2)) for x in xs])
        fl = 2 * B * HW * HW * C * O * k * k
        print(f"C{C:4d} O{O:4d} {HW:2d}x{HW:<2d} k{k}: {tn:8.1f} | {ts:8.1f} | {tl:8.1f} | {fl / tn / 1e6:7.1f} | {by / tn / 1e3:7.0f}")
    gemm_table()
    print("channel kernels: [P,C] bf16: us reduce2(stats) GB/s | us affine_act GB/s")
    for (P, C) in [(131072, 64), (131072, 256), (32768, 512), (8192, 1024), (2048, 2048)]:
        nb = max(2, min(8, int(200e6 // (P * C * 2)) + 1))
        ys = [torch.randn(P, C, device="cuda").bfloat16() for _ in range(nb)]
        outs = [torch.empty_like(y) for y in ys]
        sums = torch.zeros(2 * C, device="cuda")
        coef = torch.randn(3 * C, device="cuda")
        t1 = graph_time([(lambda y=y: call("fdt_chan_reduce2", y, y, None, None, sums, i64(P), C, 0)) for y in ys])
        t2 = graph_time([(lambda y=y, o=o: call("fdt_chan_affine_act", y, None, None, o, coef, i64(P), C, 1)) for y, o in zip(ys, outs)])
        print(f"P{P:7d} C{C:5d}: {t1:7.1f} us {P * C * 2 / t1 / 1e3:6.0f} GB/s | {t2:7.1f} us {P * C * 4 / t2 / 1e3:6.0f} GB/s")


if __name__ == "__main__":
    main()
