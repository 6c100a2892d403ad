"""Data-parallel engine.

Replaces ``torch.nn.parallel.DistributedDataParallel(model, device_ids=[rank])``
(``resnet50_test.py:716``).  Design, B200-first:

* gradients live in ONE flat fp32 buffer (:class:`FlatParameters`) in NVLink symmetric memory;
* **bucketed reduce-scatter overlapped with backward** (the reference's DDP reducer overlaps 25 MB
  bucket all-reduces with backward): the model marks its stage boundaries
  (:func:`fdt.ops.bucket_boundary`); when backward crosses one, the gradients of the stage behind it are
  final and their reduce-scatter (``csrc/comm.cu::reduce_kernel<1,*>``: in-switch ``multimem.ld_reduce``
  + 1/world + Σg²) starts on a communication stream while the rest of backward runs;
* **sharded optimizer step** (ZeRO-1/2): the reference runs ``NGD.step`` on every rank over identical
  averaged gradients (``ngd_optimizer.py:452-508``).  Here each rank owns whole tensors, receives only
  *their* mean gradients, preconditions / updates only them (1/world of the NGD work and state), and
* **all-gather by stage overlapped with the next forward**: the updated fp32 parameters are pulled
  from their owners stage by stage (``all_gather_range_kernel``, which also emits the bf16 operand
  shadow); stage *s+1* arrives while stage *s* computes (:func:`fdt.ops.stage_begin`);
* parameters are broadcast from rank 0 once at construction as one message (N2);
* BatchNorm buffers stay per-rank (the reference re-broadcasts 21 small buffers every forward
  through ``broadcast_buffers=True``; its own statistics are per-rank anyway — no SyncBN);
* ``state_dict`` keeps the ``module.`` key prefix the reference's checkpoints have (§5.4).

``FDT_DDP_SHARD=0`` (or ``shard_optimizer=False``) restores the replicated form: one fused all-reduce
kernel after backward and the full optimizer step on every rank.  Without symmetric memory (gloo on
CPU, NCCL-only boxes) the sharded step runs on library ``reduce_scatter`` / ``all_gather`` calls.
"""
from __future__ import annotations

import os
from typing import List, Optional

import torch
import torch.nn as nn

from .comm import Communicator
from .flat import FlatParameters
from . import dist as pdist


def _blocks_for(elems: int, cap: int) -> int:
    b = (elems // 4 + 1023) // 1024
    return max(1, min(cap, b))


class DataParallelEngine(nn.Module):
    def __init__(self, module: nn.Module, comm: Optional[Communicator] = None,
                 flat: Optional[FlatParameters] = None, allocator=None, shard_optimizer: Optional[bool] = None,
                 overlap: Optional[bool] = None):
        super().__init__()
        self.module = module
        self.comm = comm or Communicator()
        dev = next(module.parameters()).device
        fused_pred = None
        symm = self.comm.symm
        if shard_optimizer is None:
            shard_optimizer = os.environ.get("FDT_DDP_SHARD", "1") == "1"
        self.sharded = bool(shard_optimizer) and self.comm.world > 1 and flat is None
        self._fused_end = 0
        self._tail_ctx = None
        if self.sharded:
            self._init_sharded(module, dev, allocator, overlap)
            return
        want_fused = (symm is not None and allocator is not None and self.comm.world > 1
                      and os.environ.get("FDT_FUSED_WGRAD", "0") == "1")
        if want_fused:
            fused_pred = lambda p: bool(getattr(p, "_fdt_wgrad_fused", False))
        self.flat = flat or FlatParameters(list(module.parameters()), dev, allocator=allocator,
                                           world_size=self.comm.world, fused_first=fused_pred)
        self.comm.broadcast_(self.flat.param_flat, src=0)
        for b in module.buffers():
            self.comm.broadcast_(b, src=0)
        self._stale = True
        if want_fused and symm.owns(self.flat.grad_flat) and symm.multicast_addr(self.flat.grad_flat) and self.flat.fused_end:
            self._enable_fused_wgrad()

    def _enable_fused_wgrad(self):
        """Opt-in (``FDT_FUSED_WGRAD=1``; measured 2×B200: 16.1 ms/step vs 14.5 ms with the stand-alone
        fused all-reduce, so it is off by default — multimem.red sustains ~315 GB/s per GPU, below the
        ~550-700 GB/s of the two-shot kernel).
        Gradients of GEMM-shaped parameters are all-reduced INSIDE the wgrad kernels: every finished
        fp32 tile is scaled by 1/world and added into all ranks' gradient buffers with multimem.red
        (in-switch reduction), overlapping the transfer with the rest of backward tile by tile.  Only
        the small "tail" of the buffer (BN affine, biases, layers on the library path) still goes
        through the stand-alone fused all-reduce."""
        symm, flat = self.comm.symm, self.flat
        mc = symm.multicast_addr(flat.grad_flat)
        alpha = 1.0 / self.comm.world
        for g in flat.groups:
            if g.offset >= flat.fused_end:
                continue
            for i, p in enumerate(g.params):
                p._fdt_mc_grad = (mc + 4 * (g.offset + i * g.numel), alpha)
        self._fused_end = flat.fused_end
        n_tail = flat.total - flat.fused_end
        self._tail_ctx = symm.sub_ctx(flat.grad_flat, flat.fused_end) if n_tail > 0 else None

    # ------------------------------------------------------------------ sharded-step form
    def _init_sharded(self, module, dev, allocator, overlap):
        comm = self.comm
        symm = comm.symm
        self._native = symm is not None and allocator is not None and dev.type == "cuda"
        if overlap is None:
            overlap = os.environ.get("FDT_DDP_OVERLAP", "1") == "1"
        bucket_of = getattr(module, "grad_bucket_of", None) if (self._native and overlap) else None
        self.flat = FlatParameters(list(module.parameters()), dev, allocator=allocator if self._native else None,
                                   param_allocator=allocator if self._native else None, world_size=comm.world,
                                   shard=True, bucket_of=bucket_of, equal_regions=not self._native)
        self.region = self.flat.local_region(comm.rank)
        comm.broadcast_(self.flat.param_flat, src=0)
        for b in module.buffers():
            comm.broadcast_(b, src=0)
        self._gathered, self._stale = True, True
        self._fused_sq = None
        self._launched: List[bool] = []
        self._stage_events = None
        if not self._native:
            return
        flat = self.flat
        self._comm_stream = torch.cuda.Stream(dev)
        nb = flat.num_buckets
        cap_rs = int(os.environ.get("FDT_COMM_BLOCKS", "64"))
        self._rs_blocks = [_blocks_for(max(hi - lo for lo, hi in flat.bucket_ranges[k]), min(cap_rs, 64)) for k in range(nb)]
        # forward stage s <-> gradient bucket nb-1-s: the same sub-range of every rank's region
        # A range far larger than its peers' (one rank owns the 62 MB token embedding) is taken out of the pull all-gather —
        # world-1 ranks pulling it at once would be limited by that one owner's egress — and broadcast in two phases.
        self._ag_tables, self._ag_blocks, self._ag_bcast = [], [], []
        for s in range(nb):
            rng = list(flat.bucket_ranges[nb - 1 - s])
            mean = sum(hi - lo for lo, hi in rng) / comm.world
            big = []
            if os.environ.get("FDT_BCAST_2PHASE", "1") == "1":
                for r, (lo, hi) in enumerate(rng):
                    if hi - lo > max(1 << 20, 2 * mean):
                        big.append((r, lo, hi, _blocks_for(hi - lo, 85)))
                        rng[r] = (lo, lo)
            self._ag_bcast.append(big)
            self._ag_tables.append(torch.tensor([v for lo_hi in rng for v in lo_hi], dtype=torch.int64, device=dev))
            self._ag_blocks.append(_blocks_for(max(hi - lo for lo, hi in rng), 128))
        if flat.shadow_flat is None:
            flat.shadow_flat = torch.empty(flat.total, dtype=torch.bfloat16, device=dev)
        self._launched = [False] * nb
        if nb > 1:
            module._fdt_bucket_cb = self._on_bucket_ready
            module._fdt_stage_cb = self._on_stage_begin

    @property
    def optimizer_state_sharded(self) -> bool:
        return self.sharded

    def _launch_bucket(self, k: int):
        """Reduce-scatter of gradient bucket ``k`` on the communication stream, after everything enqueued so far."""
        if self._launched[k]:
            return
        self._launched[k] = True
        main = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(main)
        self._comm_stream.wait_event(ev)
        lo, hi = self.flat.bucket_ranges[k][self.comm.rank]
        with torch.cuda.stream(self._comm_stream):
            self.comm.symm.reduce_scatter_range_(self.flat.grad_flat, lo, hi, self._rs_blocks[k], channel=1)

    def _on_bucket_ready(self, k: int):
        # buckets must go out in the same order on every rank: k-th boundary crossed -> buckets 0..k
        for j in range(k + 1):
            self._launch_bucket(j)

    def _on_stage_begin(self, s: int):
        evs = self._stage_events
        if evs is None:
            return
        torch.cuda.current_stream().wait_event(evs[s])
        if s == len(evs) - 1:
            self._stage_events = None

    def _gather_native(self):
        """All-gather of the updated parameters (+ bf16 shadow), stage by stage on the communication stream; the compute
        stream waits per stage (``stage_begin``), so stage s+1 is in flight while stage s computes."""
        flat, symm = self.flat, self.comm.symm
        main = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(main)
        self._comm_stream.wait_event(ev)
        evs = []
        with torch.cuda.stream(self._comm_stream):
            n = len(self._ag_tables)
            for s in range(n):
                symm.all_gather_range_(flat.param_flat, self._ag_tables[s], flat.shadow_flat, self._ag_blocks[s], channel=1)
                for owner, lo, hi, blocks in self._ag_bcast[s]:
                    symm.bcast_range_(flat.param_flat, owner, lo, hi, flat.shadow_flat, blocks, channel=1)
                if s == n - 1:
                    flat.refresh_dgrad_shadow()          # backward's transposed operands need the whole shadow
                e = torch.cuda.Event()
                e.record(self._comm_stream)
                evs.append(e)
        if getattr(self.module, "_fdt_stage_cb", None) is not None and torch.is_grad_enabled() and self.module.training:
            self._stage_events = evs                     # the model's forward waits stage by stage
        else:
            main.wait_event(evs[-1])

    def _pre_forward_sharded(self):
        if not self._gathered:
            if self._native:
                self._gather_native()
                self._gathered, self._stale = True, False
                return
            self.comm.all_gather_(self.flat.param_flat, self.region)
            self._gathered, self._stale = True, True
        if self._stale:
            self.flat.refresh_shadow()
            self._stale = False

    def _reduce_sharded(self):
        self._fused_sq = None
        if not self._native:
            self.comm.reduce_scatter_mean_(self.flat.grad_flat, self.region)
            return
        if self._stage_events is not None:               # a forward that did not visit every stage
            torch.cuda.current_stream().wait_event(self._stage_events[-1])
            self._stage_events = None
        for k in range(len(self._launched)):
            self._launch_bucket(k)
        with torch.cuda.stream(self._comm_stream):
            self._fused_sq = self.comm.symm.take_sq()    # global Σg² accumulated by the bucket kernels
            done = torch.cuda.Event()
            done.record(self._comm_stream)
        torch.cuda.current_stream().wait_event(done)
        self._launched = [False] * len(self._launched)

    # ------------------------------------------------------------------ common interface
    def forward(self, *a, **kw):
        self.pre_forward()
        return self.module(*a, **kw)

    def optimizer_params(self):
        if self.sharded:
            return self.flat.owned_params(self.comm.rank)
        return self.flat.params

    def pre_forward(self):
        """Make the parameters (and their bf16 operand shadow) current: gather what the owners updated (sharded step) or
        re-cast the local buffer (replicated step)."""
        if self.sharded:
            return self._pre_forward_sharded()
        if self._stale:
            self.flat.refresh_shadow()
            self._stale = False

    def post_step(self):
        self._stale = True
        if self.sharded:
            self._gathered = False

    def reduce_gradients(self):
        if self.sharded:
            return self._reduce_sharded()
        self._fused_sq = None
        if self._fused_end:
            symm, flat = self.comm.symm, self.flat
            if self._tail_ctx is not None:      # its entry barrier also orders the peers' multimem.red traffic
                sq_tail = symm.all_reduce_mean_ctx_(self._tail_ctx, flat.total - self._fused_end)
            else:
                symm.barrier_on_stream(flat.grad_flat)
                sq_tail = 0.0
            from .. import ops
            self._fused_sq = ops._cu().sqnorm(flat.grad_flat[: self._fused_end]) + sq_tail
            return
        self.comm.all_reduce_mean_(self.flat.grad_flat)
        symm = self.comm.symm
        if symm is not None and symm.owns(self.flat.grad_flat) and self.comm.world > 1:
            self._fused_sq = symm.last_sqnorm         # Σg² of the averaged gradient, from the same kernel

    def local_grad(self) -> torch.Tensor:
        if self.sharded:
            lo, hi = self.region
            return self.flat.grad_flat[lo:hi]
        return self.flat.grad_flat

    def grad_sq_norm(self) -> torch.Tensor:
        if getattr(self, "_fused_sq", None) is not None:
            return self._fused_sq
        if self.sharded:
            g = self.local_grad()
            from .. import ops
            sq = (ops._cu().sqnorm(g) if (g.is_cuda and ops.use_native(g)) else torch.dot(g, g)).reshape(1)
            return self.comm.all_reduce_sum_(sq)[0]
        g = self.flat.grad_flat
        if g.is_cuda:
            from .. import ops
            if ops.use_native(g):
                return ops._cu().sqnorm(g)
        return torch.dot(g, g)

    def load_state_dict(self, *a, **kw):
        self._stale = True
        self.flat.shadow_fresh = False          # parameters change behind the optimizer's back: re-cast the whole shadow
        return super().load_state_dict(*a, **kw)

    def zero_grad(self):
        self.flat.zero_grad()
        if self._fused_end:
            # nobody may add into this rank's buffer before it is cleared (and nobody clears while a
            # slower peer still reads last step's gradients)
            self.comm.symm.barrier_on_stream(self.flat.grad_flat)

    def full_state_dict(self):
        if self.sharded:
            with torch.no_grad():
                self.pre_forward()           # collective: every rank ends up with every owner's current parameters
        return self.state_dict()


DDP = DataParallelEngine
