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
if what in ("conv1x1", "conv1x1_stats", "conv3x3"):
    C, O, HW, k = (128, 128, 16, 3) if what == "conv3x3" else (64, 256, 32, 1)
    x = torch.randn(B, C, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    w = (torch.randn(O, C, k, k, device="cuda") / math.sqrt(C * k * k)).bfloat16().contiguous(memory_format=torch.channels_last)
    sums = torch.zeros(2 * O, device="cuda") if what == "conv1x1_stats" else None     # BN statistics in the epilogue
    for _ in range(5):
        y = g.conv_fwd(x, w, 1, k // 2, sums)
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
elif what == "gemm2":
    a = torch.randn(8192, 8192, device="cuda").bfloat16()
    b = torch.randn(8192, 8192, device="cuda").bfloat16()
    for _ in range(4):
        y = g.gemm2_nt(a, b)
elif what == "gemm1":
    a = torch.randn(8192, 8192, device="cuda").bfloat16()
    b = torch.randn(8192, 8192, device="cuda").bfloat16()
    for _ in range(4):
        y = torch.empty(8192, 8192, dtype=torch.bfloat16, device="cuda")
        call("fdt_gemm_nt_ex", a, b, y, None, None, None, None, 8192, 8192, 8192, 0)
elif what == "conv_s2":
    C, O, HW, k = 128, 128, 32, 3
    x = torch.randn(B, C, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    w = (torch.randn(O, C, k, k, device="cuda") / math.sqrt(C * k * k)).bfloat16().contiguous(memory_format=torch.channels_last)
    dy = torch.randn(B, O, HW // 2, HW // 2, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    for _ in range(4):
        y = g.conv_fwd(x, w, 2, 1)
        dx = g.conv_dgrad(dy, w, (B, C, HW, HW), 2, 1)
elif what == "attn512":
    from fdt.ops import attention_sm100 as at
    Bt, H, L = 16, 8, 512
    qkv = torch.randn(Bt, L, 3 * H * 64, device="cuda").bfloat16().requires_grad_(True)
    for _ in range(3):
        o = at.attention_packed(qkv, H, None, 0.1, False)
        o.backward(torch.ones_like(o))
elif what == "rowwise":
    from fdt import ops
    x = torch.randn(64, 128, 512, device="cuda", requires_grad=True)
    a = torch.ones(512, device="cuda", requires_grad=True)
    b = torch.zeros(512, device="cuda", requires_grad=True)
    tok = torch.randn(30522, 512, device="cuda", requires_grad=True)
    pos = torch.randn(512, 512, device="cuda", requires_grad=True)
    seg = torch.randn(3, 512, device="cuda", requires_grad=True)
    ids = torch.randint(0, 30522, (64, 128), device="cuda")
    logits = torch.randn(128, 10, device="cuda", requires_grad=True)
    ya, yb = torch.randint(0, 10, (128,), device="cuda"), torch.randint(0, 10, (128,), device="cuda")
    for _ in range(3):
        y = ops.layernorm_ref(x, a, b)
        y.backward(torch.ones_like(y))
        e = ops.embed(ids, torch.zeros_like(ids), tok, pos, seg, 22.6)
        e.backward(torch.ones_like(e))
        l = ops.mixup_cross_entropy(logits, ya, yb, torch.tensor(0.3, device="cuda"))
        l.backward()
        z = ops.head_mix(torch.randn(64, 512, device="cuda").bfloat16(), torch.randperm(64, device="cuda"),
                         torch.tensor(0.3, device="cuda"), 0.1, True)
        c = ops.cls_gather(x)
elif what == "optim":
    n = 23_500_000
    p_ = torch.randn(n, device="cuda"); g_ = torch.randn(n, device="cuda"); buf = torch.zeros(n, device="cuda")
    v_ = torch.zeros(n, device="cuda"); s_ = torch.zeros(n, device="cuda"); x0 = p_.clone()
    hyper = torch.tensor([0.01, 0.9, 0.0], device="cuda")
    hy2 = torch.tensor([0.01, 0.0, 0.9], device="cuda")
    sq = torch.zeros((), device="cuda"); skipped = torch.zeros((), device="cuda")
    sh = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    for _ in range(3):
        sq.zero_()
        call("fdt_sqnorm", g_, i64(n), sq)
        call("fdt_clip_wd", g_, p_, i64(n), sq, 10.0, 1e-4, skipped)
        call("fdt_sgd_momentum_dev", p_, g_, buf, i64(n), hyper, 0, 0, sh)
        call("fdt_madgrad_dev", p_, g_, v_, s_, x0, i64(n), hy2, 0.0, 1e-6, 0, 1)
        call("fdt_cast_bf16", p_, sh, i64(n))
elif what == "input_head":
    from fdt import ops
    from fdt.ops import cuda_ops as cu
    x8 = torch.randint(0, 255, (128, 32, 32, 3), dtype=torch.uint8, device="cuda")
    perm = torch.randperm(128, device="cuda")
    feat = torch.randn(128, 2048, 4, 4, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    fcw = torch.randn(10, 2048, device="cuda", requires_grad=True); fcb = torch.zeros(10, device="cuda", requires_grad=True)
    theta = torch.randn(128, 1, 1, 1, device="cuda", requires_grad=True)
    for _ in range(3):
        o = cu.input_aug_mixup(x8, True, perm, torch.tensor(0.4, device="cuda"), (0.49, 0.48, 0.45), (0.25, 0.24, 0.26))
        m = cu.input_aug_meta_mixup(x8, theta, perm, (0.49, 0.48, 0.45), (0.25, 0.24, 0.26))
        m.backward(torch.ones_like(m))
        lg = ops.pool_fc(feat, fcw, fcb)
        lg.sum().backward()
elif what == "ngd_small":
    from fdt.optim import ngd as M
    fb = M.FactorBatch(1, 27, 64, 4.0, -1, 4, 0.1, torch.float32, "cuda")
    for it in range(4):
        X = torch.randn(1, 64, 27, device="cuda")
        M.precondition(fb, X, left=True)
torch.cuda.synchronize()
print("ok", what)
