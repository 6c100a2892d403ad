"""Mixup family: input mixup, "meta" mixup (per-sample λ), attention-map mixup, and the
matching loss / accuracy helpers.

Capability parity with ``resnet50_test.py:355-457`` (and the copies in
``transformer_test.py:186-202``), re-designed so that nothing crosses to the host:
λ ~ Beta(α, α) is drawn on the device (the reference samples on the CPU and calls ``.item()``,
``resnet50_test.py:359``), the permutation is a device ``randperm``, and the blend for CUDA
tensors is the fused augmentation + gather + lerp kernel in ``csrc/input_head.cu`` (``fdt_input_aug_mixup*``).

Differences from the reference that are deliberate: :class:`MetaMixup` is a
*persistent* module whose per-sample λ is a real learnable parameter — what the reference's
README claims — while ``compat=True`` reproduces what the reference code actually does
(a fresh ``sigmoid(U(0,1))`` every call, ``resnet50_test.py:525``).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from .. import ops


def _native_draws(device, generator) -> bool:
    """One-kernel device-side draws (csrc/input_head.cu::fdt_mixup_sample) when no torch generator pins the stream."""
    return generator is None and torch.device(device).type == "cuda" and not ops._FORCE_TORCH


def sample_mixup(alpha: float, n: int, device, generator: Optional[torch.Generator] = None):
    """(λ ~ Beta(α, α), random permutation of n) — on CUDA one tiny kernel for both, otherwise the torch draws."""
    if _native_draws(device, generator):
        return ops._cu().mixup_sample(alpha, n, torch.device(device))
    return sample_beta(alpha, device, generator=generator), randperm(n, device, generator)


def sample_beta(alpha: float, device, shape=(), generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """Draw Beta(α, α) on ``device`` without a host sync (two Gamma draws)."""
    if shape == () and _native_draws(device, generator):
        return ops._cu().mixup_sample(alpha, None, torch.device(device))[0]
    if alpha <= 0:
        return torch.full(shape, float(alpha), device=device)
    a = torch.full((2,) + tuple(shape), float(alpha), device=device)
    g = torch._standard_gamma(a, generator=generator) if generator is not None else torch._standard_gamma(a)
    return g[0] / (g[0] + g[1]).clamp_min(1e-20)


def randperm(n: int, device, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    if _native_draws(device, generator) and n > 0:
        return ops._cu().mixup_sample(None, n, torch.device(device))[1]
    return torch.randperm(n, device=device, generator=generator)


def blend(x: torch.Tensor, index: torch.Tensor, lam: torch.Tensor) -> torch.Tensor:
    """lam * x + (1 - lam) * x[index]; ``lam`` is 0-d or broadcastable per-sample."""
    if ops.use_native(x):
        return ops._cu().mixup_blend(x, index, lam)
    lam = lam.to(x.dtype)
    if lam.dim() == 1:
        lam = lam.view(-1, *([1] * (x.dim() - 1)))
    return lam * x + (1.0 - lam) * x[index]


def mixup_data(x, y, alpha=0.99, intra_only=False, generator=None):
    """Returns ``(mixed_x, y_a, y_b, lam)``; ``lam`` is a 0-d device tensor.

    ``intra_only`` leaves same-class pairs unmixed (resnet50_test.py:364-371; vectorised here
    instead of a per-sample Python loop)."""
    lam, index = sample_mixup(alpha, x.size(0), x.device, generator)
    y_a, y_b = y, y[index]
    if intra_only:
        per = torch.where(y_a == y_b, torch.ones_like(lam), lam).expand(x.size(0))
        mixed = blend(x, index, per.contiguous())
    else:
        mixed = blend(x, index, lam)
    return mixed, y_a, y_b, lam


class MetaMixup(nn.Module):
    """Per-sample mixing weight λ_i = sigmoid(θ_i), θ ∈ R^{B×1×1×1} (resnet50_test.py:388-401)."""

    def __init__(self, batch_size: int, compat: bool = False, device=None):
        super().__init__()
        self.batch_size = batch_size
        self.compat = compat
        self.lam = nn.Parameter(torch.rand(batch_size, 1, 1, 1, device=device))

    def forward(self, x, y, generator=None):
        B = x.size(0)
        if self.compat:
            with torch.no_grad():
                self.lam.copy_(torch.rand(self.lam.shape, device=self.lam.device, generator=generator))
        index = randperm(B, x.device, generator)
        lam = torch.sigmoid(self.lam[:B])
        if lam.requires_grad and torch.is_grad_enabled():
            # learnable path: keep the blend differentiable w.r.t. λ
            l = lam.to(x.dtype)
            mixed = l * x + (1.0 - l) * x[index]
        else:
            mixed = blend(x, index, lam.reshape(-1))
        return mixed, y, y[index], lam


    def fused_from_uint8(self, images_u8, y, mean, std, generator=None):
        """The whole input path of ``--meta_learning`` on a raw uint8 NHWC batch in one native kernel (crop + flip +
        normalise both partners + blend with sigmoid(θ_i)); differentiable w.r.t. θ.  Returns what :meth:`forward` does."""
        B = images_u8.size(0)
        if self.compat:
            with torch.no_grad():
                self.lam.copy_(torch.rand(self.lam.shape, device=self.lam.device, generator=generator))
        index = randperm(B, images_u8.device, generator)
        theta = self.lam[:B].detach() if self.compat else self.lam[:B]     # compat: lambda is resampled, never learned
        mixed = ops._cu().input_aug_meta_mixup(images_u8, theta, index, mean, std)
        return mixed, y, y[index], torch.sigmoid(theta)


class AttnMixup(nn.Module):
    """Per-pixel λ map (B, C, W, H) with ``lam_scale_i = ||λ_i||²`` (resnet50_test.py:404-424)."""

    def __init__(self, batch_size, width, height, channel=3, device=None):
        super().__init__()
        self.lam = nn.Parameter(torch.rand(batch_size, channel, width, height, device=device))

    def forward(self, x, y, generator=None):
        index = randperm(x.size(0), x.device, generator)
        lam = torch.sigmoid(self.lam[: x.size(0)]).to(x.dtype)
        mixed = lam * x + (1.0 - lam) * x[index]
        lam_scale = lam.flatten(1).pow(2).sum(dim=1)
        return mixed, y, y[index], lam_scale


def mixup_criterion(pred, y_a, y_b, lam):
    """λ·CE(pred, y_a) + (1-λ)·CE(pred, y_b), scalar or per-sample λ, one fused kernel."""
    return ops.mixup_cross_entropy(pred, y_a, y_b, lam)


mixup_criterion_meta = mixup_criterion  # the per-sample form reduces to the same fused op


def mixup_correct(pred, y_a, y_b, lam):
    return ops.mixup_correct(pred, y_a, y_b, lam)
