"""Loss scaling for fp16 runs.

bf16 is the default compute dtype on B200 and needs no loss scaling.  For ``--dtype fp16``
parity with the reference's ``GradScaler`` / ``ShardedGradScaler`` flow
(``scale → backward → unscale_ → clip → step → update``, ``resnet50_test.py:541-548``) this
scaler keeps the scale and the growth tracker on the device: ``unscale_`` multiplies the flat
gradient buffer by 1/scale, the non-finite check is part of the trainer's fused norm/clip pass,
and ``update`` adjusts the scale with tensor ops — no ``.item()``.
"""
from __future__ import annotations

import torch


class DynamicLossScaler:
    def __init__(self, device, init_scale=2.0 ** 16, growth_factor=2.0, backoff_factor=0.5, growth_interval=2000):
        self._scale = torch.full((), float(init_scale), device=device)
        self._good = torch.zeros((), device=device)
        self.growth_factor, self.backoff_factor, self.growth_interval = growth_factor, backoff_factor, growth_interval

    def scale(self, loss):
        return loss * self._scale

    def unscale_(self, flat_grad: torch.Tensor):
        flat_grad.mul_(1.0 / self._scale)

    def update(self, found_inf: torch.Tensor):
        """``found_inf``: 0-d device tensor (1.0 when the step's gradients were non-finite)."""
        # in place (copy_) so that a captured CUDA graph keeps reading / writing the same two scalars
        bad = found_inf > 0
        good = torch.where(bad, torch.zeros_like(self._good), self._good + 1)
        grow = good >= self.growth_interval
        self._scale.copy_(torch.where(bad, self._scale * self.backoff_factor,
                                      torch.where(grow, self._scale * self.growth_factor, self._scale)))
        self._good.copy_(torch.where(grow, torch.zeros_like(good), good))

    def get_scale(self) -> float:
        return float(self._scale)
