"""Training-step engine (the hot loops of ``resnet50_test.py:521-567`` and
``transformer_test.py:241-283``).

One :class:`Trainer` step is::

    zero flat grads → forward + loss (bf16 on CUDA) → backward → engine.reduce_gradients()
    → global-norm clip (device scalar, non-finite guard) → optimizer.step() → engine.post_step()

with **no host synchronisation anywhere** (the reference syncs for ``loss.item()``, the
GradScaler's found-inf check, every NGD factor and four times for its progress line).  Because
of that the whole step can be captured in a CUDA graph: ``use_cuda_graph=True`` captures one
graph per optimizer step kind ("update" = Fisher refresh, "plain") after an eager warm-up and
replays them, so a ResNet-50 step at 32×32 — hundreds of small kernels — costs one launch.

bf16 replaces the reference's fp16 autocast + GradScaler (``resnet50_test.py:533-548``): no loss
scaling is needed; ``fdt.engine.amp.DynamicLossScaler`` provides the GradScaler-equivalent
(device-side, sync-free) for ``--dtype fp16`` runs.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, Optional, Tuple

import os

import torch
import torch.nn as nn

from ..parallel.ddp import DataParallelEngine
from ..parallel.comm import Communicator


def as_engine(model: nn.Module) -> nn.Module:
    """Wrap a bare module so that it has the flat-buffer engine interface."""
    if hasattr(model, "reduce_gradients"):
        return model
    return DataParallelEngine(model, Communicator())



class _nvtx:
    """NVTX range around a phase of the step when FDT_NVTX=1 (for nsys / ncu --nvtx timelines);
    a no-op otherwise and on CPU."""
    on = os.environ.get("FDT_NVTX", "0") == "1"

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if _nvtx.on and torch.cuda.is_available():
            torch.cuda.nvtx.range_push(self.name)

    def __exit__(self, *exc):
        if _nvtx.on and torch.cuda.is_available():
            torch.cuda.nvtx.range_pop()
        return False


class Trainer:
    def __init__(self, engine: nn.Module, optimizer, forward_loss: Callable, clip_norm: Optional[float] = 10.0,
                 use_cuda_graph: bool = False, scaler=None, graph_warmup: int = 11):
        self.engine = as_engine(engine)
        self.optimizer = optimizer
        self.forward_loss = forward_loss
        self.clip_norm = clip_norm
        self.scaler = scaler
        self.device = self.engine.flat.device
        self.use_graph = bool(use_cuda_graph) and self.device.type == "cuda"
        self.graph_warmup = graph_warmup
        self._graphs: Dict[str, Any] = {}
        self._static_in = None
        self._static_out = None
        self._steps = 0
        self.skipped = torch.zeros((), device=self.device)   # steps whose gradients were non-finite
        if self.use_graph:
            # Only optimizers that keep their schedule-dependent scalars (lr, momentum, step count) on the device can
            # be replayed: a captured graph bakes every Python float into its kernels' arguments.
            if hasattr(optimizer, "set_capturable"):
                optimizer.set_capturable(True)
            else:
                import warnings
                warnings.warn(f"{type(optimizer).__name__} is not CUDA-graph capturable (its lr / momentum / step count "
                              f"are host scalars that a graph replay would freeze); running the step eagerly")
                self.use_graph = False

    # ------------------------------------------------------------------ one eager step
    def _local_grad(self) -> torch.Tensor:
        return self.engine.local_grad()

    def _clip(self):
        g = self._local_grad()
        sq = self.engine.grad_sq_norm()
        fused = getattr(self.optimizer, "fused_clip_wd", None)
        if fused is not None and g.is_cuda and fused(sq, self.clip_norm, self.skipped):
            return                          # clip + non-finite guard + weight decay ran as one pass over the gradients
        if g.is_cuda:
            from .. import ops
            if ops.use_native(g):       # one fused kernel: clip coefficient, non-finite guard, scaling
                ops._cu().clip_scale_(g, sq, self.clip_norm, self.skipped)
                return
        finite = torch.isfinite(sq)
        if self.clip_norm is not None:
            coef = (self.clip_norm / (sq.sqrt() + 1e-6)).clamp(max=1.0)
        else:
            coef = torch.ones_like(sq)
        coef = torch.where(finite, coef, torch.zeros_like(coef))
        g.mul_(coef)
        torch.nan_to_num_(g, nan=0.0, posinf=0.0, neginf=0.0)
        self.skipped += (~finite).float()

    def _eager_step(self, batch) -> Tuple[torch.Tensor, Any]:
        eng = self.engine
        if self.device.type == "cuda":
            from ..ops import cuda_ops
            cuda_ops.begin_step_seeds(self.device)           # one add: fresh Philox counters for every random kernel of the step
        with _nvtx("fdt/gather+zero"):
            eng.zero_grad()
            eng.pre_forward()
        with _nvtx("fdt/forward"):
            loss, aux = self.forward_loss(eng, batch)
        with _nvtx("fdt/backward"):
            if self.scaler is not None:
                self.scaler.scale(loss).backward()
            else:
                loss.backward()
            if self.device.type == "cuda":
                from ..ops import cuda_ops
                cuda_ops.join_side_streams()              # side-stream weight gradients have landed
        with _nvtx("fdt/reduce+clip"):
            eng.reduce_gradients()
            if self.scaler is not None:
                self.scaler.unscale_(self._local_grad())
                skipped_before = self.skipped.clone()
            self._clip()
            if self.scaler is not None:
                self.scaler.update(self.skipped - skipped_before)     # device-side found-inf → scale back-off / growth
        with _nvtx("fdt/optimizer"):
            self.optimizer.step()
            if hasattr(eng, "post_step"):
                eng.post_step()
        return loss.detach(), aux

    # ------------------------------------------------------------------ graph capture
    def _kind(self) -> str:
        fn = getattr(self.optimizer, "next_step_kind", None)
        return fn() if fn is not None else "plain"

    def _copy_in(self, batch):
        if self._static_in is None:
            self._static_in = tuple(b.clone() if torch.is_tensor(b) else b for b in batch)
        else:
            for s, b in zip(self._static_in, batch):
                if torch.is_tensor(s):
                    s.copy_(b, non_blocking=True)
        return self._static_in

    def _prime_constants(self, batch):
        """Hook for anything that must exist before capture (device constants are created lazily
        during the eager warm-up steps, so nothing to do by default)."""
        return None

    def step(self, batch) -> Tuple[torch.Tensor, Any]:
        """``batch`` is a tuple of device tensors.  Returns ``(loss, aux)`` (device tensors; with
        graphs they are static buffers overwritten by the next step)."""
        self._steps += 1
        if hasattr(self.optimizer, "sync_lr"):
            self.optimizer.sync_lr()
        if not self.use_graph:
            return self._eager_step(batch)
        kind = self._kind()
        if kind == "warm" or self._steps <= self.graph_warmup:
            out = self._eager_step(batch)
            if self._steps == self.graph_warmup:
                self._prime_constants(batch)
            return out
        if self._static_in is not None and any(
                torch.is_tensor(s_) and torch.is_tensor(b_) and s_.shape != b_.shape for s_, b_ in zip(self._static_in, batch)):
            return self._eager_step(batch)       # odd-sized batch (e.g. a ragged last one): graphs are shape-specific
        safe = getattr(self.optimizer, "graph_safe", None)
        if safe is not None and not self._graphs and not safe():
            import warnings
            warnings.warn(f"{type(self.optimizer).__name__} is not on its graph-capturable path for these parameters; "
                          f"running the step eagerly")
            self.use_graph = False
            return self._eager_step(batch)
        static = self._copy_in(batch)
        from ..ops import _lib
        if kind not in self._graphs:
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            n0 = _lib.LAUNCHES.count
            with torch.cuda.graph(g):
                out = self._eager_step(static)
            n_native = _lib.LAUNCHES.count - n0                         # native launches recorded in this graph
            self._graphs[kind] = (g, out, n_native)
            # capture only records: run the step that was just captured, so this batch is not dropped and `out` holds
            # real values (the optimizer's Python-side counters already advanced while capturing)
            g.replay()
            return out
        g, out, n_native = self._graphs[kind]
        g.replay()
        _lib.LAUNCHES.add("graph_replay", n_native)                      # they run again on every replay
        adv = getattr(self.optimizer, "advance_without_python", None)
        if adv is not None:
            adv()
        return out
