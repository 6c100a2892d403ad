"""Pure-PyTorch definitions of every fused op (the oracle).

These are plain differentiable compositions of ``torch`` primitives.  They serve three
purposes: (1) the CPU execution path (``--device cpu``; the reference has none,
``resnet50_test.py:727``), (2) the numerical oracle the sm_100a kernels are tested
against, (3) the specification of each op's exact semantics, including the reference's
quirks that matter for checkpoint parity:

* conv + *batch-stat* normalisation with unbiased variance and eps added to the standard
  deviation, no affine, no running stats, same maths in eval (``resnet.py:75-100``);
* LayerNorm with unbiased std and eps outside the square root (``transformer.py:239-242``);
* attention mask compat mode (fill value ``-1e-9``, ``transformer.py:189``);
* two-target mixup cross-entropy (``resnet50_test.py:451-457``).
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F

CELU_ALPHA = 0.075  # resnet.py:240


def activation(x: torch.Tensor, act: str) -> torch.Tensor:
    if act == "none":
        return x
    if act == "relu":
        return F.relu(x)
    if act == "celu":
        return F.celu(x, alpha=CELU_ALPHA)
    if act == "gelu":
        return F.gelu(x)
    if act == "tanh":
        return torch.tanh(x)
    raise ValueError(f"unknown activation {act!r}")


def batchstat_norm(y: torch.Tensor, eps: float) -> torch.Tensor:
    """(y - mean) / (sqrt(var_unbiased) + eps) over (N, H, W) — resnet.py:85-100."""
    mean = y.mean(dim=(0, 2, 3), keepdim=True)
    std = y.var(dim=(0, 2, 3), unbiased=True, keepdim=True).sqrt()
    return (y - mean) / (std + eps)


def conv_bnstat_act(x, weight, stride=1, padding=0, eps=1e-3, act="none", residual=None):
    """conv → batch-stat normalise → (+ residual) → activation (FusedConvBN semantics)."""
    y = F.conv2d(x, weight, stride=stride, padding=padding)
    out = batchstat_norm(y, eps)
    if residual is not None:
        out = out + residual
    return activation(out, act)


def conv_bnaffine_act(x, weight, gamma, beta, running_mean, running_var, num_batches_tracked,
                      stride=1, padding=0, eps=1e-5, momentum=0.1, training=True,
                      act="none", residual=None):
    """conv → ``nn.BatchNorm2d`` maths (biased var, eps inside sqrt, affine, running stats)."""
    y = F.conv2d(x, weight, stride=stride, padding=padding)
    if training and num_batches_tracked is not None:
        num_batches_tracked.add_(1)
    out = F.batch_norm(y, running_mean, running_var, gamma, beta, training, momentum, eps)
    if residual is not None:
        out = out + residual
    return activation(out, act)


def pool_fc(x, weight, bias):
    """AdaptiveAvgPool2d(1) + flatten + Linear — resnet.py:279-281."""
    return F.linear(x.mean(dim=(2, 3)), weight, bias)


def mixup_cross_entropy(logits, y_a, y_b, lam):
    """mean_i( lam_i * CE(logits_i, y_a_i) + (1 - lam_i) * CE(logits_i, y_b_i) ).

    ``lam`` may be a python float, a 0-d tensor (plain mixup, resnet50_test.py:451-452) or a
    per-sample tensor (meta mixup, :455-457).  Both reference forms reduce to this mean.
    """
    logp = F.log_softmax(logits.float(), dim=-1)
    la = -logp.gather(1, y_a.view(-1, 1)).squeeze(1)
    lb = -logp.gather(1, y_b.view(-1, 1)).squeeze(1)
    if not torch.is_tensor(lam):
        lam = torch.tensor(float(lam), device=logits.device)
    lam = lam.reshape(-1).float()
    return (lam * la + (1.0 - lam) * lb).mean()


def mixup_correct(logits, y_a, y_b, lam):
    """λ-weighted match count used for the training accuracy (resnet50_test.py:553-558)."""
    pred = logits.argmax(dim=1)
    if not torch.is_tensor(lam):
        lam = torch.tensor(float(lam), device=logits.device)
    lam = lam.reshape(-1).float()
    return (lam * pred.eq(y_a).float() + (1.0 - lam) * pred.eq(y_b).float()).sum()


def layernorm_ref(x, a_2, b_2, eps=1e-6):
    """a * (x - mean) / (std_unbiased + eps) + b — transformer.py:239-242."""
    mean = x.mean(-1, keepdim=True)
    std = x.std(-1, keepdim=True)
    return a_2 * (x - mean) / (std + eps) + b_2


def linear(x, weight, bias=None, act="none"):
    return activation(F.linear(x, weight, bias), act)


def attention(q, k, v, key_mask=None, dropout_p=0.0, training=True, mask_compat=False):
    """softmax(q kᵀ / sqrt(d)) v with key-padding mask; q, k, v: [B, h, L, d].

    ``key_mask`` is [B, L] with 1 = keep.  ``mask_compat=True`` reproduces the reference's
    fill value of ``-1e-9`` (transformer.py:189), which leaves padding effectively unmasked.
    """
    d = q.size(-1)
    scores = torch.matmul(q, k.transpose(-2, -1)) / math.sqrt(d)
    if key_mask is not None:
        m = key_mask[:, None, None, :] == 0
        scores = scores.masked_fill(m, -1e-9 if mask_compat else float("-inf"))
    p = F.softmax(scores.float(), dim=-1).to(q.dtype)
    if dropout_p > 0.0 and training:
        p = F.dropout(p, dropout_p, True)
    return torch.matmul(p, v)


def embed(tokens, token_types, tok_w, pos_w, seg_w, scale):
    """(pos[0:L] + tok[tokens] + seg[types]) * sqrt(d_model) — transformer.py:150-156."""
    L = tokens.size(1)
    return (pos_w[:L].unsqueeze(0) + tok_w[tokens] + seg_w[token_types]) * scale


def mlp_head(x, W1, b1, W2, b2, bias_grad_mean_compat=False):
    """Linear → ReLU → Linear with (1, H)-shaped biases (transformer.py:292-371).

    ``bias_grad_mean_compat`` reproduces the reference's bias gradients (mean over the batch
    instead of sum, transformer.py:311,327).
    """
    if bias_grad_mean_compat:
        B = x.shape[0]
        b1 = _ScaleGrad.apply(b1, 1.0 / B)
        b2 = _ScaleGrad.apply(b2, 1.0 / B)
    h = F.relu(F.linear(x, W1, b1.reshape(-1)))
    return F.linear(h, W2, b2.reshape(-1))


class _ScaleGrad(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, s):
        ctx.s = s
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return g * ctx.s, None


def dropout_add(x, residual, p, training):
    """residual + dropout(x) (sub-layer connections, transformer.py:254-259, 271-275)."""
    if p > 0.0 and training:
        x = F.dropout(x, p, True)
    return x + residual
