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
    def forward(ctx, x, weight, bias, act, drop_p=0.0):
        from .cuda_ops import ACT, _next_seed_off
        from ._lib import call, i64
        shp = x.shape
        x16 = x.to(BF16).reshape(-1, shp[-1]).contiguous()
        w16 = _w16(weight).contiguous()
        b32 = None if bias is None else bias.detach().float().reshape(-1).contiguous()
        so = None
        if act == "gelu" or drop_p > 0:
            # keep the pre-activation for the backward; activation + dropout are one pass, the mask is regenerated
            # from (seed, offset) in the backward instead of being stored
            z = g.gemm_nt(x16, w16, b32, "none")
            out = torch.empty_like(z)
            so = _next_seed_off(z.device) if drop_p > 0 else None
            call("fdt_act_dropout_fwd", z, out, i64(z.numel()), ACT[act], float(drop_p), so)
            saved = z if act == "gelu" else out
        else:
            out = g.gemm_nt(x16, w16, b32, act)
            saved = out
        ctx.save_for_backward(x16, w16, saved, weight, bias, so)
        ctx.act, ctx.shp, ctx.xdtype, ctx.drop_p = act, shp, x.dtype, float(drop_p)
        ctx.bshape = None if bias is None else bias.shape
        return out.view(*shp[:-1], w16.shape[0])

    @staticmethod
    def backward(ctx, gout):
        from .cuda_ops import ACT, _flat_grad_slot, _wdgrad
        from ._lib import call, i64
        x16, w16, saved, weight, bias, so = ctx.saved_tensors
        drop_p = ctx.drop_p
        N = w16.shape[0]
        g2 = gout.reshape(-1, N)
        if g2.dtype not in (BF16, torch.float32):
            g2 = g2.float()
        if not g2.is_contiguous():
            g2 = g2.contiguous()
        T = g2.shape[0]
        dx = dw = db = None
        # one pass: gz = gout ∘ act'(saved) in bf16 (the GEMM operand) and db += Σ_rows gz
        need_db = ctx.bshape is not None and ctx.needs_input_grad[2]
        db_slot = _flat_grad_slot(bias) if need_db else None
        if need_db and db_slot is None:
            db = torch.zeros(N, dtype=torch.float32, device=g2.device)
        reuse = ctx.act == "none" and drop_p == 0 and g2.dtype == BF16   # the incoming gradient already is the operand
        gz = g2 if reuse else torch.empty(T, N, dtype=BF16, device=g2.device)
        if need_db or not reuse:
            args = (g2, int(g2.dtype == torch.float32), saved if ctx.act != "none" else None, None if reuse else gz,
                    (db_slot if db_slot is not None else db) if need_db else None, i64(T), N, ACT[ctx.act])
            if drop_p > 0:
                call("fdt_linear_bwd_prep_dropout", *args, drop_p, so)
            else:
                call("fdt_linear_bwd_prep", *args)
        if db is not None:
            db = db.view(ctx.bshape)
        if ctx.needs_input_grad[0]:
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
                slot = _flat_grad_slot(weight)       # accumulate straight into the flat gradient buffer
                dw = wgrad_sm100.linear_wgrad(gz, x16, slot) if slot is not None else wgrad_sm100.linear_wgrad(gz, x16)
                if dw is True:
                    dw = None
                elif dw is None:
                    dw = (gz.t() @ x16).float()
        return dx, dw, db, None, None


def linear(x, weight, bias, act, drop_p: float = 0.0) -> Optional[torch.Tensor]:
    K = x.shape[-1]
    N = weight.shape[0]
    if K % 64 or N % 64 or not g.available():
        return None
    if drop_p > 0 and act not in ("none", "gelu"):
        return None          # act' is taken from the (dropped, rescaled) output for relu / tanh: not supported fused
    return _Linear.apply(x, weight, bias, act, drop_p)
