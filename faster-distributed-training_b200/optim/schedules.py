"""Learning-rate schedules used by the reference drivers.

* ResNet + NGD      → ``MultiStepLR([10, 20], gamma=0.2)``          (resnet50_test.py:489)
* ResNet + MADGRAD  → ``CosineAnnealingLR(T_max=200)``              (resnet50_test.py:494)
* Transformer       → ``OneCycleLR(max_lr=5*lr, epochs, steps_per_epoch)`` stepped once per
  epoch (transformer_test.py:224-226, 291 — SURVEY.md Q16; ``per_step=True`` steps it properly)
* tuning            → ``StepLR(2, gamma)`` / ``MultiStepLR([10, 15], 0.1)``

``scale_lr`` reproduces the reference's ``lr * 4`` under ``--distributed`` (hard-coded for 4
GPUs, Q5); by default the factor is the actual world size.
"""
from __future__ import annotations

from torch.optim import lr_scheduler


def scale_lr(lr: float, distributed: bool, world_size: int, compat_factor: float | None = None) -> float:
    if not distributed:
        return lr
    return lr * (compat_factor if compat_factor is not None else world_size)


def build_scheduler(kind: str, optimizer, **kw):
    if kind == "multistep":
        return lr_scheduler.MultiStepLR(optimizer, kw.get("milestones", [10, 20]), gamma=kw.get("gamma", 0.2))
    if kind == "cosine":
        return lr_scheduler.CosineAnnealingLR(optimizer, T_max=kw.get("T_max", 200))
    if kind == "step":
        return lr_scheduler.StepLR(optimizer, kw.get("step_size", 2), gamma=kw.get("gamma", 0.75))
    if kind == "onecycle":
        # cycle_momentum needs a 'momentum' (or 'betas') entry in the param groups, which NGD/MADGRAD have
        return lr_scheduler.OneCycleLR(optimizer, max_lr=kw["max_lr"], epochs=kw["epochs"],
                                       steps_per_epoch=kw["steps_per_epoch"],
                                       cycle_momentum=kw.get("cycle_momentum", True))
    raise ValueError(f"unknown scheduler {kind!r}")
