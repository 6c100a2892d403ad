"""Collective front end.

``Communicator`` is what the DDP / FSDP engines call.  It picks, per call:

* **native NVLink path** — CUDA tensors that live in the symmetric heap
  (:mod:`fdt.parallel.symm`): the hand-written sm_100a kernels in ``csrc/comm.cu`` read and
  write peer memory directly (P2P ``ld/st`` over NVSwitch, or ``multimem`` when a multicast
  mapping exists) and fuse the mean scaling, the squared-norm partial for gradient clipping and
  the bf16 shadow cast into the same pass;
* **library path** — ``torch.distributed`` (NCCL on GPUs, gloo on CPU): the correctness oracle,
  the timed baseline and the CPU plumbing mode.

Scalar metrics are packed into one small tensor and reduced once (N4: the reference issues three
separate 4-byte all-reduces per epoch, ``resnet50_test.py:616-619``).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist

from . import dist as pdist


class Communicator:
    def __init__(self, group=None, symm=None):
        self.group = group
        self.symm = symm  # fdt.parallel.symm.SymmetricHeap or None
        self.world = pdist.world_size()
        self.rank = pdist.rank()

    # ------------------------------------------------------------------ gradients
    def all_reduce_mean_(self, flat: torch.Tensor) -> torch.Tensor:
        if self.world == 1:
            return flat
        if self.symm is not None and self.symm.owns(flat):
            return self.symm.all_reduce_mean_(flat)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
        return flat.div_(self.world)

    def reduce_scatter_mean_(self, flat: torch.Tensor, region: Tuple[int, int]) -> torch.Tensor:
        """After the call ``flat[region]`` holds the mean over ranks of that slice."""
        lo, hi = region
        if self.world == 1:
            return flat[lo:hi]
        if self.symm is not None and self.symm.owns(flat):
            return self.symm.reduce_scatter_mean_(flat, lo, hi)
        out = flat[lo:hi]
        if flat.is_cuda:
            dist.reduce_scatter_tensor(out, flat, op=dist.ReduceOp.SUM, group=self.group)
        else:  # gloo has no reduce_scatter: all-reduce then keep the local slice
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
        return out.div_(self.world)

    def all_gather_(self, flat: torch.Tensor, region: Tuple[int, int]) -> torch.Tensor:
        """Every rank contributes ``flat[region]``; afterwards ``flat`` is complete everywhere."""
        if self.world == 1:
            return flat
        lo, hi = region
        if self.symm is not None and self.symm.owns(flat):
            return self.symm.all_gather_(flat, lo, hi)
        if flat.is_cuda:
            dist.all_gather_into_tensor(flat, flat[lo:hi], group=self.group)
        else:
            chunks = list(flat.view(self.world, -1).unbind(0))
            dist.all_gather(chunks, flat[lo:hi].clone(), group=self.group)
        return flat

    def broadcast_(self, t: torch.Tensor, src: int = 0) -> torch.Tensor:
        if self.world > 1:
            dist.broadcast(t, src=src, group=self.group)
        return t

    # ------------------------------------------------------------------ scalars
    def all_reduce_sum_(self, packed: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=self.group)
        return packed

    def all_reduce_max_(self, packed: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(packed, op=dist.ReduceOp.MAX, group=self.group)
        return packed

    def barrier(self):
        if self.world > 1:
            dist.barrier(group=self.group)
