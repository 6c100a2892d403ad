"""MADGRAD and MirrorMADGRAD — the reference's non-``--ngd`` default optimizers.

The reference imports them from the external ``madgrad`` pip package
(``resnet50_test.py:32,493``; ``transformer_test.py:34,220``), which is not part of its tree and
not installed on this image, so they are re-implemented here from the published algorithm
(Defazio & Jelassi, "Adaptivity without Compromise", 2021): dual averaging of gradients with a
cube-root second-moment denominator and ``lr * sqrt(k + 1)`` step weighting.  On CPU (and for parameters
outside a flat buffer) the state is per tensor; when the parameters live in a
:class:`fdt.parallel.flat.FlatParameters` buffer on a GPU the whole update is ONE launch of the fused
kernel in ``csrc/optim_multi.cu`` over the flat range, with ``lr``, the step count ``k`` and the
momentum in device memory — so the step can be captured in a CUDA graph and still follow LR /
momentum schedules (``set_capturable`` / ``sync_lr``, the interface :class:`fdt.engine.Trainer`
looks for).
"""
from __future__ import annotations

import math

import torch
from torch.optim import Optimizer


class _FlatMadgradMixin:
    """Flat-buffer, device-scalar execution path shared by MADGRAD and MirrorMADGRAD."""
    _mirror = False

    def _flat_init(self):
        self._flat = {}          # gi -> dict(span, v, s, x0, hyper, host)
        self.capturable = False

    def set_capturable(self, flag: bool = True):
        self.capturable = bool(flag)

    def next_step_kind(self) -> str:
        return "plain" if self._flat else "warm"      # the first (eager) step builds the flat state

    def graph_safe(self) -> bool:
        """True when every param group runs on the flat device-scalar path (the per-tensor path keeps lr / k on the host)."""
        return len(self._flat) == len(self.param_groups) and all(bool(st) for st in self._flat.values())

    def _hyper_values(self, group, k):
        return (float(group["lr"]) + float(group["eps"]), float(k), float(group["momentum"]))

    def sync_lr(self):
        for gi, group in enumerate(self.param_groups):
            st = self._flat.get(gi)
            if not st:
                continue
            vals = (float(group["lr"]) + float(group["eps"]), float(group["momentum"]))
            if st["host"] != vals:
                if torch.cuda.is_current_stream_capturing():
                    continue
                h = st["hyper"]
                h[0].fill_(vals[0])
                h[2].fill_(vals[1])
                st["host"] = vals

    def _flat_state(self, gi, group):
        st = self._flat.get(gi)
        if st is not None:
            return st or None
        from ..parallel.flat import FlatParameters
        params = [p for p in group["params"] if p.grad is not None]
        span = FlatParameters.common_span(params) if params and params[0].is_cuda else None
        if span is not None:
            from .. import ops
            if not ops.use_native(params[0]):
                span = None
        if span is None:
            self._flat[gi] = {}
            return None
        flat, lo, hi = span
        P = flat.param_flat[lo:hi]
        st = {"span": span, "v": torch.zeros_like(P), "s": None if self._mirror else torch.zeros_like(P),
              "x0": P.detach().clone(),
              "hyper": torch.tensor(self._hyper_values(group, 0), dtype=torch.float32, device=P.device),
              "host": (float(group["lr"]) + float(group["eps"]), float(group["momentum"]))}
        self._flat[gi] = st
        return st

    def _flat_step(self, gi, group) -> bool:
        st = self._flat_state(gi, group)
        if st is None:
            return False
        from .. import ops
        flat, lo, hi = st["span"]
        ops._cu().madgrad_dev_(flat.param_flat[lo:hi], flat.grad_flat[lo:hi], st["v"], st["s"], st["x0"], st["hyper"],
                               group["weight_decay"], group["eps"], self._mirror)
        return True

    def state_dict(self):
        sd = super().state_dict()
        sd["fdt_flat"] = {gi: {k: (v.detach().clone() if torch.is_tensor(v) else None) for k, v in st.items()
                               if k in ("v", "s", "x0", "hyper")} for gi, st in self._flat.items() if st}
        return sd

    def load_state_dict(self, state):
        state = dict(state)
        flat = state.pop("fdt_flat", {})
        for gi, saved in flat.items():
            group = self.param_groups[int(gi)]
            for p in group["params"]:
                if p.grad is None:
                    p.grad = torch.zeros_like(p)
            st = self._flat_state(int(gi), group)
            if st is None or saved["v"].numel() != st["v"].numel():
                raise ValueError("flat MADGRAD state does not match this rank's parameter span")
        super().load_state_dict(state)
        for gi, saved in flat.items():
            st = self._flat[int(gi)]
            for k in ("v", "s", "x0", "hyper"):
                if saved.get(k) is not None and st[k] is not None:
                    st[k].copy_(saved[k])


class MADGRAD(_FlatMadgradMixin, Optimizer):
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
        self._flat_init()

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        self.sync_lr()
        for gi, group in enumerate(self.param_groups):
            if self._flat_step(gi, group):
                continue
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


class MirrorMADGRAD(_FlatMadgradMixin, Optimizer):
    _mirror = True

    """Mirror-descent form: the iterate ``z`` is updated directly in the changing cube-root metric,
    ``z ← z − λ_k g_k / (v_{k+1}^{1/3} + eps)``, and ``p`` is its momentum average."""

    def __init__(self, params, lr=1e-2, momentum=0.9, weight_decay=0.0, eps=1e-6):
        if momentum < 0 or momentum >= 1:
            raise ValueError(f"Momentum {momentum} must be in the range [0,1)")
        if lr <= 0:
            raise ValueError(f"Learning rate {lr} must be positive")
        super().__init__(params, dict(lr=lr, eps=eps, momentum=momentum, weight_decay=weight_decay))
        self._flat_init()

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        self.sync_lr()
        for gi, group in enumerate(self.param_groups):
            if self._flat_step(gi, group):
                continue
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
