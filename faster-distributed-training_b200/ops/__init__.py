"""Fused-op front end.

Every model and optimizer in the framework calls the functions below.  On CPU tensors they
run the pure-PyTorch oracle in :mod:`fdt.ops.reference`; on CUDA tensors they run the
hand-written sm_100a kernels from ``csrc/`` (loaded by :mod:`fdt.ops._lib`) through the
autograd wrappers in :mod:`fdt.ops.cuda_ops`.  If the native library is missing the call raises
(``FDT_FORCE_TORCH=1`` explicitly selects the oracle; the numerics tests use it to produce
reference values on the same device).  A shape the native kernels do not cover runs on the
library path (cuDNN / cuBLAS / ATen) for that call only — never silently: every such call is
counted per op in ``fdt.ops._lib.FALLBACKS``, the totals are printed at exit, and
``FDT_STRICT=1`` turns the first one into a ``NativeCoverageError``.  The default training
configurations of both models (ResNet-50 at 32x32, the 6-layer transformer up to L=512) take
no fallback; ``tests/test_strict_native.py`` asserts that under ``FDT_STRICT=1``.
"""
from __future__ import annotations

import os

import torch

from . import reference as ref

_FORCE_TORCH = os.environ.get("FDT_FORCE_TORCH", "0") == "1"


def force_torch(flag: bool) -> None:
    """Route CUDA tensors through the PyTorch oracle (tests / debugging only)."""
    global _FORCE_TORCH
    _FORCE_TORCH = bool(flag)


def use_native(t: torch.Tensor) -> bool:
    return t.is_cuda and not _FORCE_TORCH


def _cu():
    from . import cuda_ops
    return cuda_ops


def launch_counter():
    """Counter of native kernel launches issued by this process (see ``fdt.ops._lib``)."""
    from . import _lib
    return _lib.LAUNCHES


def conv_bnstat_act(x, weight, stride=1, padding=0, eps=1e-3, act="none", residual=None, arena=None, passthrough=False):
    """``arena``: optional ``(BNArena, offset)`` — a statistics slot cleared once per forward pass by the model.
    ``passthrough``: also return the input as a second output (an alias); whoever else consumes the input should consume
    that alias, so its gradient is added inside this unit's data-gradient GEMM (see models/resnet.py::_Block)."""
    if use_native(x):
        return _cu().conv_bnstat_act(x, weight, stride, padding, eps, act, residual, arena, passthrough)
    out = ref.conv_bnstat_act(x, weight, stride, padding, eps, act, residual)
    return (out, x) if passthrough else out


def conv_bnaffine_act(x, weight, gamma, beta, running_mean, running_var, num_batches_tracked,
                      stride=1, padding=0, eps=1e-5, momentum=0.1, training=True,
                      act="none", residual=None, arena=None, passthrough=False):
    if use_native(x):
        return _cu().conv_bnaffine_act(x, weight, gamma, beta, running_mean, running_var,
                                       num_batches_tracked, stride, padding, eps, momentum,
                                       training, act, residual, arena, passthrough)
    out = ref.conv_bnaffine_act(x, weight, gamma, beta, running_mean, running_var,
                                num_batches_tracked, stride, padding, eps, momentum,
                                training, act, residual)
    return (out, x) if passthrough else out


def pool_fc(x, weight, bias):
    if use_native(x):
        return _cu().pool_fc(x, weight, bias)
    return ref.pool_fc(x, weight, bias)


def mixup_cross_entropy(logits, y_a, y_b, lam):
    if use_native(logits):
        return _cu().mixup_cross_entropy(logits, y_a, y_b, lam)
    return ref.mixup_cross_entropy(logits, y_a, y_b, lam)


def mixup_correct(logits, y_a, y_b, lam):
    return ref.mixup_correct(logits, y_a, y_b, lam)


def layernorm_ref(x, a_2, b_2, eps=1e-6, passthrough=False):
    """``passthrough``: also return the input (an alias) for the residual branch to consume — see cuda_ops._LayerNorm."""
    if use_native(x):
        return _cu().layernorm_ref(x, a_2, b_2, eps, passthrough)
    y = ref.layernorm_ref(x, a_2, b_2, eps)
    return (y, x) if passthrough else y


def linear(x, weight, bias=None, act="none", dropout_p=0.0, training=False):
    """act(x Wᵀ + b) [→ dropout(dropout_p) when ``training``]."""
    p = float(dropout_p) if training else 0.0
    if use_native(x):
        return _cu().linear(x, weight, bias, act, p)
    out = ref.linear(x, weight, bias, act)
    return torch.nn.functional.dropout(out, p, True) if p > 0 else out


def attention(q, k, v, key_mask=None, dropout_p=0.0, training=True, mask_compat=False):
    if use_native(q):
        return _cu().attention(q, k, v, key_mask, dropout_p, training, mask_compat)
    return ref.attention(q, k, v, key_mask, dropout_p, training, mask_compat)


def qkv_attention(x, weights, biases, heads, key_mask=None, dropout_p=0.0, training=True, mask_compat=False):
    """Fused Q/K/V projection (one GEMM over the back-to-back weights) + attention on the packed result.  Returns
    [B, L, d_model], or None when the fused path does not apply (CPU, parameters not adjacent in a flat buffer) — the
    caller then runs the three projections separately."""
    if use_native(x):
        return _cu().qkv_attention(x, weights, biases, heads, key_mask, dropout_p if training else 0.0, mask_compat)
    return None


def embed(tokens, token_types, tok_w, pos_w, seg_w, scale):
    if use_native(tok_w):
        return _cu().embed(tokens, token_types, tok_w, pos_w, seg_w, scale)
    return ref.embed(tokens, token_types, tok_w, pos_w, seg_w, scale)


def mlp_head(x, W1, b1, W2, b2, bias_grad_mean_compat=False):
    if use_native(x):
        return _cu().mlp_head(x, W1, b1, W2, b2, bias_grad_mean_compat)
    return ref.mlp_head(x, W1, b1, W2, b2, bias_grad_mean_compat)


def linear_dropout_add(x, weight, bias, residual, p, training):
    """``residual + dropout(x Wᵀ + b, p)`` — the closing Linear of a sub-layer with its dropout and residual add."""
    p = float(p) if training else 0.0
    if use_native(x):
        out = _cu().linear_dropout_add(x, weight, bias, residual, p)
        if out is not None:
            return out
    return dropout_add(linear(x, weight, bias), residual, p, True)


def cls_gather(x):
    """First-token rows of the [B, L, D] residual stream (the pooler's input, transformer.py:99-101)."""
    if use_native(x):
        return _cu().cls_gather(x)
    return x[:, 0, :].contiguous()


def head_mix(z, perm, lam, p, training):
    """Pooled-vector dropout followed by manifold mixup ``lam*z + (1-lam)*z[perm]`` (transformer.py:70-80); ``perm`` None
    = no mixing."""
    p = float(p) if training else 0.0
    if use_native(z):
        return _cu().head_mix(z, perm, lam, p)
    z = torch.nn.functional.dropout(z, p, True) if p > 0 else z
    if perm is None:
        return z
    l = lam.to(z.dtype)
    return l * z + (1.0 - l) * z[perm]


def dropout_add(x, residual, p, training):
    if use_native(x):
        return _cu().dropout_add(x, residual, p, training)
    return ref.dropout_add(x, residual, p, training)


class _BucketBoundary(torch.autograd.Function):
    """Identity whose backward tells the parallel engine that a phase of the backward pass is over."""

    @staticmethod
    def forward(ctx, x, cb, k):
        ctx.cb, ctx.k = cb, k
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        ctx.cb(ctx.k)
        return g, None, None


def bucket_boundary(x, owner, k: int):
    """Mark the point of a model's forward pass after which only parameters of gradient buckets < ``k``... are used:
    when backward crosses it, every kernel producing gradients of bucket ``k`` has been enqueued, and the engine
    wrapping ``owner`` (it installs ``owner._fdt_bucket_cb``) may start that bucket's reduce-scatter on its
    communication stream while the rest of backward runs.  A no-op without an engine (or without grad)."""
    cb = getattr(owner, "_fdt_bucket_cb", None)
    if cb is None or not torch.is_grad_enabled() or not x.requires_grad:
        return x
    return _BucketBoundary.apply(x, cb, k)


def stage_begin(owner, k: int):
    """Forward-pass counterpart: parameters of stage ``k`` are about to be used (the engine may still be gathering them on
    its communication stream; ``owner._fdt_stage_cb`` makes the compute stream wait for that gather)."""
    cb = getattr(owner, "_fdt_stage_cb", None)
    if cb is not None:
        cb(k)
