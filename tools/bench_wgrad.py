#!/usr/bin/env python
"""Micro-benchmark: tcgen05 split-K weight-gradient kernel vs cuDNN / cuBLAS on the ResNet-50 and transformer
shapes (graph-timed, NB rotating buffer sets so operands are not L2-resident across calls)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench_gemm import graph_time  # noqa: E402
from fdt.ops import wgrad_sm100 as wg  # noqa: E402


def main():
    torch.backends.cudnn.benchmark = True
    B = 128
    print("conv wgrad (N=128): us native | us cudnn | TFLOP/s native")
    for (C, O, HW, k) in [(64, 64, 32, 3), (64, 64, 32, 1), (64, 256, 32, 1), (256, 64, 32, 1), (128, 128, 16, 3), (128, 512, 16, 1),
                          (512, 128, 16, 1), (256, 256, 8, 3), (256, 1024, 8, 1), (1024, 256, 8, 1), (512, 512, 4, 3),
                          (512, 2048, 4, 1), (2048, 512, 4, 1)]:
        by = (B * C * HW * HW + B * O * HW * HW) * 2
        nb = max(2, min(8, int(160e6 // by) + 1))
        mk = lambda ch: torch.randn(B, ch, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
        xs = [mk(C) for _ in range(nb)]
        dys = [mk(O) for _ in range(nb)]
        dw = torch.zeros(O, k, k, C, device="cuda")
        w = torch.randn(O, C, k, k, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
        tn = graph_time([(lambda x=x, dy=dy: wg.conv_wgrad(dy, x, (O, C, k, k), 1, k // 2, dw)) for x, dy in zip(xs, dys)])
        tl = graph_time([(lambda x=x, dy=dy: torch.ops.aten.convolution_backward(
            dy, x, w, None, [1, 1], [k // 2, k // 2], [1, 1], False, [0, 0], 1, [False, True, False])) for x, dy in zip(xs, dys)])
        fl = 2 * B * HW * HW * C * O * k * k
        print(f"C{C:4d} O{O:4d} {HW:2d}x{HW:<2d} k{k}: {tn:8.1f} | {tl:8.1f} | {fl / tn / 1e6:7.1f}")
    print("linear wgrad dW[N,K] = dY[M,N]^T X[M,K]: us native | us cublas | TFLOP/s native")
    for (M, N, K) in [(8192, 512, 512), (8192, 1024, 512), (8192, 512, 1024), (8192, 2048, 512), (8192, 512, 2048)]:
        nb = 6
        xs = [torch.randn(M, K, device="cuda").bfloat16() for _ in range(nb)]
        dys = [torch.randn(M, N, device="cuda").bfloat16() for _ in range(nb)]
        dw = torch.zeros(N, K, device="cuda")
        tn = graph_time([(lambda x=x, dy=dy: wg.linear_wgrad(dy, x, dw)) for x, dy in zip(xs, dys)])
        tl = graph_time([(lambda x=x, dy=dy: dy.t() @ x) for x, dy in zip(xs, dys)])
        print(f"M{M} N{N} K{K}: {tn:8.1f} | {tl:8.1f} | {2 * M * N * K / tn / 1e6:7.1f}")


if __name__ == "__main__":
    main()
