"""MADGRAD and MirrorMADGRAD — the reference's non-``--ngd`` default optimizers.

The reference imports them from the external ``madgrad`` pip package
(``resnet50_test.py:32,493``; ``transformer_test.py:34,220``), which is not part of its tree and
not installed on this image, so they are re-implemented here from the published algorithm
(Defazio & Jelassi, "Adaptivity without Compromise", 2021): dual averaging of gradients with a
cube-root second-moment denominator and ``lr * sqrt(k + 1)`` step weighting.  All state is
per-tensor and the update is expressed with ``torch._foreach`` multi-tensor ops on CPU; on CUDA
the same update is the fused multi-tensor kernel in ``csrc/optim_multi.cu``.
"""
from __future__ import annotations

import math

import torch
from torch.optim import Optimizer


class MADGRAD(Optimizer):
    def __init__(self, params, lr=1e-2, momentum=0.9, weight_decay=0.0, eps=1e-6):
        if momentum < 0 or momentum >= 1:
            raise ValueError(f"Momentum {momentum} must be in the range [0,1)")
        if lr <= 0:
            raise ValueError(f"Learning rate {lr} must be positive")
        if weight_decay < 0:
            raise ValueError(f"Weight decay {weight_decay} must be non-negative")
        if eps < 0:
            raise ValueError("Eps must be non-negative")
        super().__init__(params, dict(lr=lr, eps=eps, momentum=momentum, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            eps, lr, decay, momentum = group["eps"], group["lr"] + group["eps"], group["weight_decay"], group["momentum"]
            ck = 1.0 - momentum
            for p in group["params"]:
                if p.grad is None:
                    continue
                grad = p.grad
                state = self.state[p]
                if len(state) == 0:
                    state["k"] = 0
                    state["grad_sum_sq"] = torch.zeros_like(p)
                    state["s"] = torch.zeros_like(p)
                    if momentum != 0:
                        state["x0"] = p.detach().clone()
                k = state["k"]
                lamb = lr * math.sqrt(k + 1)
                if decay != 0:
                    grad = grad.add(p, alpha=decay)
                v, s = state["grad_sum_sq"], state["s"]
                if momentum == 0:
                    rms = v.pow(1.0 / 3.0).add_(eps)
                    x0 = p.addcdiv(s, rms, value=1.0)
                else:
                    x0 = state["x0"]
                v.addcmul_(grad, grad, value=lamb)
                rms = v.pow(1.0 / 3.0).add_(eps)
                s.add_(grad, alpha=lamb)
                if momentum == 0:
                    p.copy_(x0.addcdiv(s, rms, value=-1.0))
                else:
                    z = x0.addcdiv(s, rms, value=-1.0)
                    p.mul_(1.0 - ck).add_(z, alpha=ck)
                state["k"] = k + 1
        return loss


class MirrorMADGRAD(Optimizer):
    """Mirror-descent form: the iterate ``z`` is updated directly in the changing cube-root metric,
    ``z ← z − λ_k g_k / (v_{k+1}^{1/3} + eps)``, and ``p`` is its momentum average."""

    def __init__(self, params, lr=1e-2, momentum=0.9, weight_decay=0.0, eps=1e-6):
        if momentum < 0 or momentum >= 1:
            raise ValueError(f"Momentum {momentum} must be in the range [0,1)")
        if lr <= 0:
            raise ValueError(f"Learning rate {lr} must be positive")
        super().__init__(params, dict(lr=lr, eps=eps, momentum=momentum, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            eps, lr, decay, momentum = group["eps"], group["lr"] + group["eps"], group["weight_decay"], group["momentum"]
            ck = 1.0 - momentum
            for p in group["params"]:
                if p.grad is None:
                    continue
                grad = p.grad
                state = self.state[p]
                if len(state) == 0:
                    state["k"] = 0
                    state["grad_sum_sq"] = torch.zeros_like(p)
                    state["z"] = p.detach().clone()
                k = state["k"]
                lamb = lr * math.sqrt(k + 1)
                if decay != 0:
                    grad = grad.add(p, alpha=decay)
                v, z = state["grad_sum_sq"], state["z"]
                v.addcmul_(grad, grad, value=lamb)
                rms = v.pow(1.0 / 3.0).add_(eps)
                z.addcdiv_(grad, rms, value=-lamb)
                if momentum == 0:
                    p.copy_(z)
                else:
                    p.mul_(1.0 - ck).add_(z, alpha=ck)
                state["k"] = k + 1
        return loss
