"""Fused attention on tcgen05 (``csrc/attention_sm100.cu``): L <= 512, d = 64, KV-tile loop with online softmax.

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


def _launch_fwd(q, k, v, o, lse, km, B, H, L, D, ldq, p_drop, mask_compat):
    so = None
    if p_drop > 0:
        from .cuda_ops import _next_seed_off
        so = _next_seed_off(o.device)
    scale = 1.0 / math.sqrt(D)
    call("fdt_attn_fwd", q, k, v, o, lse, km, B, H, L, ldq, H * D, float(scale), float(p_drop), so, int(bool(mask_compat)))
    return so, scale


def _dq_acc(B, H, L, D, device):
    """fp32 accumulation buffer for dQ when a sequence spans several 128-key tiles (cleared inside the call)."""
    return torch.empty(B * L, H * D, dtype=torch.float32, device=device) if L > 128 else None


class _Attention(torch.autograd.Function):
    """q, k, v: separate [B, h, L, d] views of [B, L, h*d] projection outputs."""

    @staticmethod
    def forward(ctx, q, k, v, key_mask, p_drop, mask_compat):
        B, H, L, D = q.shape
        q, k, v = _as_rows(q), _as_rows(k), _as_rows(v)
        ld = H * D
        o = torch.empty(B, L, ld, dtype=BF16, device=q.device)
        lse = torch.empty(B, H, L, dtype=torch.float32, device=q.device)
        km = None if key_mask is None else key_mask.to(torch.int64).contiguous()
        so, scale = _launch_fwd(q, k, v, o, lse, km, B, H, L, D, ld, p_drop, mask_compat)
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
        call("fdt_attn_bwd", q, k, v, o, go, lse, dq, dk, dv, _dq_acc(B, H, L, D, q.device), km, B, H, L, ld, ld, ld,
             float(scale), float(p_drop), so, int(bool(mask_compat)), launches=1 if L <= 128 else 3)
        f = lambda t: t.view(B, L, H, D).transpose(1, 2)
        return f(dq), f(dk), f(dv), None, None, None


class _AttentionPacked(torch.autograd.Function):
    """qkv: ONE [B, L, 3*h*d] bf16 matrix (the output of the fused Q/K/V projection GEMM); the kernels read the three
    column slices in place (row stride 3*h*d) and the backward writes dQ | dK | dV into one matrix of the same layout —
    the operand of the fused projection's data- and weight-gradient GEMMs.  Returns o [B, L, h*d]."""

    @staticmethod
    def forward(ctx, qkv, H, key_mask, p_drop, mask_compat):
        B, L, W3 = qkv.shape
        hd = W3 // 3
        D = hd // H
        qkv = qkv.contiguous()
        o = torch.empty(B, L, hd, dtype=BF16, device=qkv.device)
        lse = torch.empty(B, H, L, dtype=torch.float32, device=qkv.device)
        km = None if key_mask is None else key_mask.to(torch.int64).contiguous()
        flat = qkv.view(B * L, W3)
        so, scale = _launch_fwd(flat[:, :hd], flat[:, hd:2 * hd], flat[:, 2 * hd:], o, lse, km, B, H, L, D, W3, p_drop, mask_compat)
        ctx.save_for_backward(qkv, o, lse, km, so)
        ctx.cfg = (B, H, L, D, scale, p_drop, mask_compat)
        return o

    @staticmethod
    def backward(ctx, go):
        qkv, o, lse, km, so = ctx.saved_tensors
        B, H, L, D, scale, p_drop, mask_compat = ctx.cfg
        hd = H * D
        W3 = 3 * hd
        go = go.to(BF16).contiguous()
        dqkv = torch.empty(B, L, W3, dtype=BF16, device=qkv.device)
        flat, dflat = qkv.view(B * L, W3), dqkv.view(B * L, W3)
        call("fdt_attn_bwd", flat[:, :hd], flat[:, hd:2 * hd], flat[:, 2 * hd:], o, go, lse,
             dflat[:, :hd], dflat[:, hd:2 * hd], dflat[:, 2 * hd:], _dq_acc(B, H, L, D, qkv.device), km, B, H, L, W3, hd, W3,
             float(scale), float(p_drop), so, int(bool(mask_compat)), launches=1 if L <= 128 else 3)
        return dqkv, None, None, None, None


def attention(q, k, v, key_mask, dropout_p, mask_compat) -> Optional[torch.Tensor]:
    B, H, L, D = q.shape
    if D != 64 or L > 512 or not available():
        return None
    return _Attention.apply(q, k, v, key_mask, float(dropout_p), bool(mask_compat))


def attention_packed(qkv, heads, key_mask, dropout_p, mask_compat) -> Optional[torch.Tensor]:
    B, L, W3 = qkv.shape
    if W3 % (3 * heads) or W3 // (3 * heads) != 64 or L > 512 or qkv.dtype != BF16 or not available():
        return None
    return _AttentionPacked.apply(qkv, int(heads), key_mask, float(dropout_p), bool(mask_compat))
