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
            if wt is not None and tuple(wt.shape) != (w16.shape[1], N):
                wt = None                            # the slot holds the fused Q|K|V transpose, not this weight's alone
            dx = g.gemm_nt(gz, wt if wt is not None else w16.t().contiguous(), None, "none")
            if dx is None:
                from ._lib import note_fallback
                note_fallback("linear_bwd", "cuBLAS data-gradient GEMM")
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
                    from ._lib import note_fallback
                    note_fallback("linear_bwd", "cuBLAS weight-gradient GEMM")
                    dw = (gz.t() @ x16).float()
        return dx, dw, db, None, None


class _LinearFused(torch.autograd.Function):
    """y = x [W_0; W_1; ...]ᵀ + [b_0; b_1; ...] for k same-shaped Linear layers whose parameters sit back to back in the
    flat buffers (the Q / K / V projections): ONE forward GEMM with N = k*out, one data-gradient GEMM against the fused
    transposed weight, one weight-gradient GEMM accumulating into the k adjacent gradient slots, one bias-gradient pass.
    ``weights`` / ``biases`` are passed so autograd sees them, but every operand is addressed through the flat buffers."""

    @staticmethod
    def forward(ctx, x, flat, w_off, b_off, k, *params):
        weights, biases = params[:k], params[k:]
        N1, K = weights[0].shape
        N = k * N1
        shp = x.shape
        x16 = x.to(BF16).reshape(-1, K).contiguous()
        w16 = flat.shadow_flat[w_off:w_off + N * K].view(N, K)
        b32 = flat.param_flat[b_off:b_off + N] if b_off is not None else None
        out = g.gemm_nt(x16, w16, b32, "none")
        ctx.save_for_backward(x16, w16)
        ctx.flat, ctx.w_off, ctx.b_off, ctx.k, ctx.shp, ctx.xdtype, ctx.leader = flat, w_off, b_off, k, shp, x.dtype, weights[0]
        return out.view(*shp[:-1], N)

    @staticmethod
    def backward(ctx, gout):
        from ._lib import call, i64
        from .cuda_ops import _wdgrad
        from . import wgrad_sm100
        x16, w16 = ctx.saved_tensors
        flat, k = ctx.flat, ctx.k
        N, K = w16.shape
        g2 = gout.reshape(-1, N)
        if g2.dtype != BF16 or not g2.is_contiguous():
            g2 = g2.to(BF16).contiguous()
        T = g2.shape[0]
        if ctx.b_off is not None:         # Σ_rows straight into the k adjacent bias-gradient slots
            call("fdt_linear_bwd_prep", g2, 0, None, None, flat.grad_flat[ctx.b_off:ctx.b_off + N], i64(T), N, 0)
        dx = None
        if ctx.needs_input_grad[0]:
            wt = _wdgrad(ctx.leader, 0)   # [in, k*out]: the fused transpose made once per step
            if wt is None or tuple(wt.shape) != (K, N):
                wt = w16.t().contiguous()
            dx = g.gemm_nt(g2, wt, None, "none").view(ctx.shp).to(ctx.xdtype)
        slot = flat.grad_flat[ctx.w_off:ctx.w_off + N * K].view(N, K)
        if wgrad_sm100.linear_wgrad(g2, x16, slot) is not True:
            from ._lib import note_fallback
            note_fallback("linear_bwd", "cuBLAS weight-gradient GEMM (fused QKV)")
            slot.add_((g2.t() @ x16).float())
        return (dx, None, None, None, None) + (None,) * (2 * k)


def linear_fused(x, weights, biases) -> Optional[torch.Tensor]:
    """See :class:`_LinearFused`; None when the parameters are not adjacent in one flat buffer (caller falls back)."""
    f = getattr(weights[0], "_fdt_flat", None)
    if f is None or not g.available():
        return None
    flat = f[0]
    if flat.shadow_flat is None or weights[0].shape[1] % 64 or weights[0].shape[0] % 64:
        return None
    w_off = flat.adjacent_block(list(weights))
    b_off = None
    if biases[0] is not None:
        b_off = flat.adjacent_block(list(biases))
        if b_off is None:
            return None
    if w_off is None or (w_off * 2) % 16 or (b_off is not None and (b_off * 4) % 16):
        return None
    return _LinearFused.apply(x, flat, w_off, b_off, len(weights), *weights, *[b for b in biases if b is not None])


class _LinearDropoutAdd(torch.autograd.Function):
    """out (fp32) = residual + dropout(x Wᵀ + b, p): the closing Linear of a pre-LN sub-layer together with its dropout and
    residual add (transformer.py:254-259, 271-275).  Forward: GEMM with bias epilogue + one dropout-add pass.  Backward: ONE
    pass turns the fp32 residual-stream gradient into the masked bf16 GEMM operand AND the bias gradient (the separate
    dropout-backward and bias-reduction passes are gone), then the data- and weight-gradient GEMMs; the residual-stream
    gradient itself is handed on untouched."""

    @staticmethod
    def forward(ctx, x, weight, bias, residual, p):
        from .cuda_ops import _next_seed_off
        from ._lib import call, i64
        shp = x.shape
        x16 = x.to(BF16).reshape(-1, shp[-1]).contiguous()
        w16 = _w16(weight).contiguous()
        b32 = None if bias is None else bias.detach().float().reshape(-1).contiguous()
        z = g.gemm_nt(x16, w16, b32, "none")
        res = residual.float().contiguous()
        out = torch.empty_like(res)
        so = _next_seed_off(z.device)          # the kernel reads the pair even for p == 0
        call("fdt_dropout_add", z, 1, res, out, i64(z.numel()), float(p), so)
        ctx.save_for_backward(x16, w16, weight, bias, so)
        ctx.p, ctx.shp, ctx.xdtype = float(p), shp, x.dtype
        return out

    @staticmethod
    def backward(ctx, gout):
        from .cuda_ops import _flat_grad_slot, _wdgrad
        from ._lib import call, i64
        from . import wgrad_sm100
        x16, w16, weight, bias, so = ctx.saved_tensors
        N, K = w16.shape
        g32 = gout.float().contiguous()
        g2 = g32.view(-1, N)
        T = g2.shape[0]
        need_db = bias is not None and ctx.needs_input_grad[2]
        db_slot = _flat_grad_slot(bias) if need_db else None
        db = None
        if need_db and db_slot is None:
            db = torch.zeros(N, dtype=torch.float32, device=g2.device)
        gz = torch.empty(T, N, dtype=BF16, device=g2.device)
        dbp = (db_slot if db_slot is not None else db) if need_db else None
        if ctx.p > 0:
            call("fdt_linear_bwd_prep_dropout", g2, 1, None, gz, dbp, i64(T), N, 0, ctx.p, so)
        else:
            call("fdt_linear_bwd_prep", g2, 1, None, gz, dbp, i64(T), N, 0)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            wt = _wdgrad(weight, 0)
            if wt is not None and tuple(wt.shape) != (K, N):
                wt = None
            dx = g.gemm_nt(gz, wt if wt is not None else w16.t().contiguous(), None, "none").view(ctx.shp).to(ctx.xdtype)
        if ctx.needs_input_grad[1]:
            slot = _flat_grad_slot(weight)
            dw = wgrad_sm100.linear_wgrad(gz, x16, slot) if slot is not None else wgrad_sm100.linear_wgrad(gz, x16)
            if dw is True:
                dw = None
            elif dw is None:
                from ._lib import note_fallback
                note_fallback("linear_bwd", "cuBLAS weight-gradient GEMM")
                dw = (gz.t() @ x16).float()
        if db is not None:
            db = db.view(bias.shape)
        return dx, dw, db, (g32 if ctx.needs_input_grad[3] else None), None


def linear_dropout_add(x, weight, bias, residual, p) -> Optional[torch.Tensor]:
    K, N = x.shape[-1], weight.shape[0]
    if K % 64 or N % 64 or not g.available() or tuple(residual.shape[:-1]) != tuple(x.shape[:-1]) or residual.shape[-1] != N:
        return None
    return _LinearDropoutAdd.apply(x, weight, bias, residual, float(p))


def linear(x, weight, bias, act, drop_p: float = 0.0) -> Optional[torch.Tensor]:
    K = x.shape[-1]
    N = weight.shape[0]
    if K % 64 or N % 64 or not g.available():
        return None
    if drop_p > 0 and act not in ("none", "gelu"):
        return None          # act' is taken from the (dropped, rescaled) output for relu / tanh: not supported fused
    return _Linear.apply(x, weight, bias, act, drop_p)
