"""Linear layer on the tcgen05 GEMM: forward with fused bias + activation epilogue, data gradient
through the same kernel with the transposed weight, weight gradient through the MN-major kernel."""
from __future__ import annotations

from typing import Optional

import torch

from . import gemm_sm100 as g
from . import reference as ref

BF16 = torch.bfloat16


def _w16(p):
    from .cuda_ops import _w16 as f
    return f(p)


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, act):
        shp = x.shape
        x16 = x.to(BF16).reshape(-1, shp[-1]).contiguous()
        w16 = _w16(weight).contiguous()
        b32 = None if bias is None else bias.detach().float().reshape(-1).contiguous()
        if act == "gelu":
            z = g.gemm_nt(x16, w16, b32, "none")          # keep the pre-activation for the backward
            out = torch.nn.functional.gelu(z)
            saved = z
        else:
            out = g.gemm_nt(x16, w16, b32, act)
            saved = out
        ctx.save_for_backward(x16, w16, saved, weight)
        ctx.act, ctx.shp, ctx.xdtype = act, shp, x.dtype
        ctx.bshape = None if bias is None else bias.shape
        return out.view(*shp[:-1], w16.shape[0])

    @staticmethod
    def backward(ctx, gout):
        x16, w16, saved, weight = ctx.saved_tensors
        N = w16.shape[0]
        gz = gout.to(BF16).reshape(-1, N)
        if ctx.act == "gelu":
            zf = saved.float()
            gz = (gz.float() * (0.5 * (1 + torch.erf(zf * 0.7071067811865476)) +
                                zf * 0.3989422804014327 * torch.exp(-0.5 * zf * zf))).to(BF16)
        elif ctx.act == "relu":
            gz = gz * (saved > 0)
        elif ctx.act == "tanh":
            gz = (gz.float() * (1 - saved.float() ** 2)).to(BF16)
        gz = gz.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            from .cuda_ops import _wdgrad
            wt = _wdgrad(weight, 0)                  # [in, out], produced once per step for all layers
            dx = g.gemm_nt(gz, wt if wt is not None else w16.t().contiguous(), None, "none")
            if dx is None:
                dx = gz @ w16
            dx = dx.view(ctx.shp).to(ctx.xdtype)
        if ctx.needs_input_grad[1]:
            from . import wgrad_sm100
            mc = getattr(weight, "_fdt_mc_grad", None)
            if mc is not None:      # in-switch all-reduce fused into the wgrad epilogue
                dw = wgrad_sm100.linear_wgrad(gz, x16, mc[0], mc[1], True)
                if dw is True:
                    dw = None
                else:
                    from .cuda_ops import _mc_accumulate
                    _mc_accumulate(weight, (gz.t() @ x16).float())
                    dw = None
            else:
                dw = wgrad_sm100.linear_wgrad(gz, x16)
                if dw is None:
                    dw = (gz.t() @ x16).float()
        if ctx.bshape is not None and ctx.needs_input_grad[2]:
            db = gz.float().sum(0).view(ctx.bshape)
        return dx, dw, db, None


def linear(x, weight, bias, act) -> Optional[torch.Tensor]:
    K = x.shape[-1]
    N = weight.shape[0]
    if K % 64 or N % 64 or not g.available():
        return None
    return _Linear.apply(x, weight, bias, act)
