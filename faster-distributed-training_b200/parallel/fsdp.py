"""Fully-sharded data-parallel engine.

Replaces ``FullyShardedDataParallel(model, auto_wrap_policy=size_based_auto_wrap_policy,
cpu_offload=CPUOffload(offload_params=True))`` (``transformer_test.py:387-392``), which for this
29 M-parameter model degenerates to one flat unit sharded 1/world per rank and parked in host
memory.  B200-first re-design:

* **Sharding by whole shape groups.**  Each rank owns complete tensors (a balanced subset of the
  shape groups of :class:`FlatParameters`), so the NGD optimizer keeps its per-tensor / per-axis
  semantics on every rank.  The reference instead preconditions a flat 1-D shard (D = shard
  size, N = 1 — a different algorithm); this is the documented decision.
* **Full-shard life cycle.**  ``pre_forward`` all-gathers the parameter buffer (every rank
  contributes its region); after backward ``reduce_gradients`` reduce-scatters so each rank holds
  the *mean* gradient of its own region only; the optimizer (built over ``optimizer_params()`` =
  owned tensors) updates just that region, so optimizer state (NGD factors, momentum) is sharded
  too (ZeRO-1/2/3 in one; C10).  Regions are padded to equal length so both collectives are single
  in-place calls — or single fused NVLink kernels when the buffers sit in symmetric memory.
* **No CPU offload**: a 180 GB B200 does not need it.
* **Global gradient norm** for clipping = Σ over ranks of the local-region Σg² (the reference
  clips each shard by its local norm, N7 / Q17).
* ``full_state_dict`` is collective-safe: all ranks gather, every rank returns the full dict
  (the reference calls ``state_dict()`` on rank 0 only and hangs, Q10).
"""
from __future__ import annotations

from typing import Optional

import os

import torch
import torch.nn as nn

from .comm import Communicator
from .flat import FlatParameters



_POISON = os.environ.get("FDT_DEBUG_POISON", "0") == "1"

class FullyShardedEngine(nn.Module):
    """Dispatches to the staged engine (per-stage units, overlapped collectives) when the buffers can live in NVLink symmetric
    memory, and to :class:`WholeModelShardedEngine` (one flat unit, library collectives) otherwise."""

    def __new__(cls, module: nn.Module, comm: Optional[Communicator] = None, allocator=None, **kw):
        if cls is FullyShardedEngine:
            c = comm or Communicator()
            staged = (c.world > 1 and c.symm is not None and allocator is not None
                      and os.environ.get("FDT_FSDP_LEGACY", "0") != "1")
            return super().__new__(StagedShardedEngine if staged else WholeModelShardedEngine)
        return super().__new__(cls)


class WholeModelShardedEngine(FullyShardedEngine):
    def __init__(self, module: nn.Module, comm: Optional[Communicator] = None, allocator=None):
        nn.Module.__init__(self)
        self.module = module
        self.comm = comm or Communicator()
        dev = next(module.parameters()).device
        self.flat = FlatParameters(list(module.parameters()), dev, allocator=allocator,
                                   world_size=self.comm.world, shard=True, param_allocator=allocator)
        self.region = self.flat.local_region(self.comm.rank)
        self.comm.broadcast_(self.flat.param_flat, src=0)
        for b in module.buffers():
            self.comm.broadcast_(b, src=0)
        self._gathered = True
        self._stale = True

    @property
    def optimizer_state_sharded(self) -> bool:
        """Each rank's optimizer holds only the tensors it owns (checkpoint sidecars are per rank)."""
        return self.comm.world > 1

    def forward(self, *a, **kw):
        self.pre_forward()
        return self.module(*a, **kw)

    def optimizer_params(self):
        return self.flat.owned_params(self.comm.rank)

    def pre_forward(self):
        if not self._gathered:
            symm = self.comm.symm
            if symm is not None and symm.owns(self.flat.param_flat):
                # one kernel: pull every peer's region over NVLink and emit the bf16 operand shadow
                if self.flat.shadow_flat is None:
                    self.flat.shadow_flat = torch.empty(self.flat.total, dtype=torch.bfloat16, device=self.flat.device)
                symm.all_gather_(self.flat.param_flat, *self.region, shadow=self.flat.shadow_flat)
                self.flat.refresh_dgrad_shadow()
                self._gathered, self._stale = True, False
                return
            self.comm.all_gather_(self.flat.param_flat, self.region)
            self._gathered = True
            self._stale = True
        if self._stale:
            self.flat.refresh_shadow()
            self._stale = False

    def reduce_gradients(self):
        self._fused_sq = None
        self.comm.reduce_scatter_mean_(self.flat.grad_flat, self.region)
        symm = self.comm.symm
        if symm is not None and symm.owns(self.flat.grad_flat) and self.comm.world > 1:
            self._fused_sq = symm.last_sqnorm         # global Σg², produced by the same kernel

    def post_step(self):
        """Call after the optimizer updated the owned region: remote regions are now stale."""
        self._gathered = self.comm.world == 1
        self._stale = True
        if _POISON and self.comm.world > 1:
            # debug mode: everything this rank does not own, and the whole operand shadow, become NaN,
            # so a gather that misses a byte — or a kernel that reads a stale copy — poisons the next loss visibly
            lo, hi = self.region
            nan = float("nan")
            self.flat.param_flat[:lo].fill_(nan)
            self.flat.param_flat[hi:].fill_(nan)
            if self.flat.shadow_flat is not None:
                self.flat.shadow_flat.fill_(nan)
            if getattr(self.flat, "dgrad_flat", None) is not None:
                self.flat.dgrad_flat.fill_(nan)

    def local_grad(self) -> torch.Tensor:
        lo, hi = self.region
        return self.flat.grad_flat[lo:hi]

    def grad_sq_norm(self) -> torch.Tensor:
        if getattr(self, "_fused_sq", None) is not None:
            return self._fused_sq
        lo, hi = self.region
        g = self.flat.grad_flat[lo:hi]
        from .. import ops
        sq = (ops._cu().sqnorm(g) if ops.use_native(g) else torch.dot(g, g)).reshape(1)
        return self.comm.all_reduce_sum_(sq)[0]

    def zero_grad(self):
        self.flat.zero_grad()

    def full_state_dict(self):
        self.pre_forward()
        return self.state_dict()


from .ddp import DataParallelEngine  # noqa: E402


class StagedShardedEngine(DataParallelEngine, FullyShardedEngine):
    """FULL_SHARD on NVLink symmetric memory: every rank owns whole tensors (optimizer state, fp32 master copy and the
    authoritative values live with the owner), the model is cut into stage units (``module.STAGES`` / ``stage_of``), and

    * the parameters of stage s+1 are gathered from their owners (``all_gather_range_kernel``: P2P pulls + bf16 operand
      shadow in the same pass; oversized single-owner ranges by a two-phase broadcast) WHILE stage s computes,
    * the gradients of a stage are reduce-scattered to their owners (in-switch ``multimem.ld_reduce``) as soon as backward
      has left the stage, overlapping the rest of backward,
    * each rank runs the NGD / MADGRAD step on what it owns only.

    This is the machinery of :class:`DataParallelEngine`'s sharded step; the engine differs from DDP in name and in
    refusing to fall back to a replicated optimizer step.  (The reference wraps the whole 29 M-parameter model in ONE
    FSDP unit parked in host memory — ``transformer_test.py:387-392``.)"""

    def __init__(self, module: nn.Module, comm: Optional[Communicator] = None, allocator=None):
        DataParallelEngine.__init__(self, module, comm, allocator=allocator, shard_optimizer=True, overlap=True)


FSDP = FullyShardedEngine
