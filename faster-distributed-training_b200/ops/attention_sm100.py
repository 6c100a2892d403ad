"""Fused attention on tcgen05 (``csrc/attention_sm100.cu``): one CTA per (batch, head), L <= 128, d = 64.

q, k, v arrive as ``[B, h, L, d]`` views of the projection outputs ``[B, L, h*d]`` (that is what
``MultiheadAttention`` produces with ``view(...).transpose(1, 2)``); the kernels address those matrices
directly through TMA at column offset ``head*64``, and the result is returned as the same kind of view, so
no head transpose is ever materialised."""
from __future__ import annotations

import math
from typing import Optional

import torch

from . import _lib
from ._lib import call

BF16 = torch.bfloat16


def available() -> bool:
    try:
        return hasattr(_lib.lib(), "fdt_attn_fwd")
    except Exception:
        return False


def _as_rows(t: torch.Tensor) -> torch.Tensor:
    """[B, h, L, d] → same values with memory laid out as [B, L, h, d] (no copy if already so)."""
    B, H, L, D = t.shape
    if t.dtype == BF16 and t.stride() == (L * H * D, D, H * D, 1):
        return t
    return t.to(BF16).transpose(1, 2).contiguous().transpose(1, 2)


class _Attention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, key_mask, p_drop, mask_compat):
        B, H, L, D = q.shape
        q, k, v = _as_rows(q), _as_rows(k), _as_rows(v)
        ld = H * D
        o = torch.empty(B, L, ld, dtype=BF16, device=q.device)
        lse = torch.empty(B, H, L, dtype=torch.float32, device=q.device)
        km = None if key_mask is None else key_mask.to(torch.int64).contiguous()
        so = None
        if p_drop > 0:
            from .cuda_ops import _next_seed_off
            so = _next_seed_off(q.device)
        scale = 1.0 / math.sqrt(D)
        call("fdt_attn_fwd", q, k, v, o, lse, km, B, H, L, ld, float(scale), float(p_drop), so, int(bool(mask_compat)))
        ctx.save_for_backward(q, k, v, o, lse, km, so)
        ctx.cfg = (B, H, L, D, scale, p_drop, mask_compat)
        return o.view(B, L, H, D).transpose(1, 2)

    @staticmethod
    def backward(ctx, go):
        q, k, v, o, lse, km, so = ctx.saved_tensors
        B, H, L, D, scale, p_drop, mask_compat = ctx.cfg
        go = _as_rows(go)
        ld = H * D
        dq = torch.empty(B, L, ld, dtype=BF16, device=q.device)
        dk = torch.empty_like(dq)
        dv = torch.empty_like(dq)
        call("fdt_attn_bwd", q, k, v, o, go, lse, dq, dk, dv, km, B, H, L, ld, float(scale), float(p_drop), so,
             int(bool(mask_compat)))
        f = lambda t: t.view(B, L, H, D).transpose(1, 2)
        return f(dq), f(dk), f(dv), None, None, None


def attention(q, k, v, key_mask, dropout_p, mask_compat) -> Optional[torch.Tensor]:
    B, H, L, D = q.shape
    if D != 64 or L > 128 or not available():
        return None
    return _Attention.apply(q, k, v, key_mask, float(dropout_p), bool(mask_compat))
