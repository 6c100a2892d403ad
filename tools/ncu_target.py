#!/usr/bin/env python
"""Tiny single-kernel driver for `ncu` captures (one GPU, a handful of launches).
usage: ncu_target.py {conv1x1|conv3x3|gemm|wgrad|attn|chan_affine|chan_reduce|ngd_eig|ngd_axis|ngd_tc|lin_prep|wprep}"""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fdt.ops import gemm_sm100 as g  # noqa: E402
from fdt.ops._lib import call, i64  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "conv1x1"
B = 128
if what in ("conv1x1", "conv3x3"):
    C, O, HW, k = (64, 256, 32, 1) if what == "conv1x1" else (128, 128, 16, 3)
    x = torch.randn(B, C, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    w = (torch.randn(O, C, k, k, device="cuda") / math.sqrt(C * k * k)).bfloat16().contiguous(memory_format=torch.channels_last)
    for _ in range(5):
        y = g.conv_fwd(x, w, 1, k // 2)
elif what == "gemm":
    a = torch.randn(8192, 8192, device="cuda").bfloat16()
    b = torch.randn(8192, 8192, device="cuda").bfloat16()
    for _ in range(4):
        y = g.gemm_nt(a, b)
elif what in ("chan_affine", "chan_reduce"):
    P, C = 131072, 256
    y = torch.randn(P, C, device="cuda").bfloat16()
    o = torch.empty_like(y)
    sums = torch.zeros(2 * C, device="cuda")
    coef = torch.randn(3 * C, device="cuda")
    for _ in range(5):
        if what == "chan_affine":
            call("fdt_chan_affine_act", y, None, None, o, coef, i64(P), C, 1)
        else:
            call("fdt_chan_reduce2", y, y, None, None, sums, i64(P), C, 0)
elif what == "wgrad":
    from fdt.ops import wgrad_sm100 as wg
    C, O, HW, k = 256, 256, 8, 3
    x = torch.randn(B, C, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    dy = torch.randn(B, O, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    for _ in range(5):
        dw = wg.conv_wgrad(dy, x, (O, C, k, k), 1, 1)
elif what == "attn":
    from fdt.ops import attention_sm100 as at
    Bt, H, L = 64, 8, 128
    mk = lambda: torch.randn(Bt, L, H * 64, device="cuda").bfloat16().view(Bt, L, H, 64).transpose(1, 2).requires_grad_(True)
    q, k_, v = mk(), mk(), mk()
    for _ in range(4):
        o = at.attention(q, k_, v, None, 0.1, False)
        o.backward(torch.ones_like(o))
elif what == "ngd_eig":
    from fdt.optim import ngd as M
    fb = M.FactorBatch(148, 576, 512, 4.0, -1, 4, 0.1, torch.float32, "cuda")
    for it in range(6):
        X = torch.randn(148, 576, 512, device="cuda")
        M.precondition(fb, X, left=False)
elif what == "ngd_axis":
    from fdt.ops import cuda_ops as cu
    X = torch.randn(3, 512, 3, 3 * 512, device="cuda")          # [n, O, kh, kw*I] view of three 512x512x3x3 gradients
    W = torch.randn(3, 2, 3, device="cuda") * 0.3
    J = torch.empty(3, 2, 3, device="cuda")
    for _ in range(4):
        cu.ngd_small_axis(X, W, 512, 3, 3 * 512, J)
elif what == "ngd_tc":
    from fdt.ops import cuda_ops as cu
    X = torch.randn(3, 512, 4608, device="cuda")                 # first-axis factor of three 512x512x3x3 gradients (transposed use)
    W = torch.randn(3, 80, 512, device="cuda") * 0.05
    J = torch.empty(3, 80, 512, device="cuda")
    for _ in range(3):
        H, out, p0 = cu.ngd_project(X, W, True)
        cu.ngd_stats_j(H, X, J, True)
elif what == "lin_prep":
    g = torch.randn(8192, 1024, device="cuda").bfloat16()
    z = torch.randn(8192, 1024, device="cuda").bfloat16()
    gz = torch.empty_like(g)
    db = torch.zeros(1024, device="cuda")
    for _ in range(4):
        call("fdt_linear_bwd_prep", g, 0, z, gz, db, i64(8192), 1024, 3)
elif what == "wprep":
    from fdt.models.resnet import resnet50
    from fdt.parallel.flat import FlatParameters
    m = resnet50(10)
    flat = FlatParameters(list(m.parameters()), device=torch.device("cuda"))
    for _ in range(4):
        flat.refresh_shadow()
torch.cuda.synchronize()
print("ok", what)
