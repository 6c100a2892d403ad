"""NVLink symmetric heap + the fused in-kernel collectives of ``csrc/comm.cu``.

``torch.distributed._symmetric_memory`` is used purely as the *allocation and handle-exchange*
layer (CUDA VMM allocations, peer mappings, optional multicast object); all data movement is done
by this framework's own kernels, which read / write the peer addresses directly.

A :class:`SymmetricHeap` owns
* the flat buffers handed out by :meth:`alloc_f32` (gradient buffer for DDP, gradient + parameter
  buffers for FSDP) — every rank can address every peer's copy,
* a signal buffer for the per-block flag barriers and a scratch word that receives the global Σg²,
and exposes the three collectives the engines need.  Each is ONE kernel launch:

=========================  ========================================================================
``all_reduce_mean_``       two-shot all-reduce + 1/world scaling + Σg² (gradient-clipping norm)
``reduce_scatter_mean_``   rank-local region reduce + scaling + *global* Σg² via peer atomics
``all_gather_``            pull all regions + emit the bf16 operand shadow in the same pass
=========================  ========================================================================
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional, Tuple

import torch
import torch.distributed as dist

from ..ops import _lib

MAX_WORLD = 8


class _CommCtx(ctypes.Structure):
    _fields_ = [("buf", ctypes.c_void_p * MAX_WORLD), ("sig", ctypes.c_void_p * MAX_WORLD),
                ("scratch", ctypes.c_void_p * MAX_WORLD), ("mc", ctypes.c_void_p),
                ("rank", ctypes.c_int), ("world", ctypes.c_int), ("pad0", ctypes.c_int), ("pad1", ctypes.c_int)]


class SymmetricHeap:
    def __init__(self, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.symm_mem = symm_mem
        self.device = torch.device(device)
        self.group = group or dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        if self.world > MAX_WORLD:
            raise RuntimeError(f"symmetric heap supports up to {MAX_WORLD} ranks on one node")
        try:
            symm_mem.enable_symm_mem_for_group(self.group.group_name)
        except Exception:
            pass
        self.sig = symm_mem.empty(2 * 64 * MAX_WORLD * 2, dtype=torch.int32, device=self.device)
        self.sig.zero_()
        self.scratch = symm_mem.empty(64, dtype=torch.float32, device=self.device)
        self.scratch.zero_()
        torch.cuda.synchronize(self.device)
        self.sig_hdl = symm_mem.rendezvous(self.sig, self.group)
        self.scratch_hdl = symm_mem.rendezvous(self.scratch, self.group)
        self.bufs: Dict[int, Tuple[torch.Tensor, object, _CommCtx]] = {}
        self.use_multicast = True
        self.last_sqnorm: Optional[torch.Tensor] = None
        dist.barrier(self.group)

    @classmethod
    def create(cls, device, group=None) -> "SymmetricHeap":
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() < 2:
            raise RuntimeError("symmetric heap needs an initialised process group with world size > 1")
        if torch.device(device).type != "cuda":
            raise RuntimeError("symmetric heap needs CUDA")
        return cls(device, group)

    # ------------------------------------------------------------------ allocation
    def alloc_f32(self, numel: int) -> torch.Tensor:
        t = self.symm_mem.empty(numel, dtype=torch.float32, device=self.device)
        t.zero_()
        torch.cuda.synchronize(self.device)
        hdl = self.symm_mem.rendezvous(t, self.group)
        ctx = _CommCtx()
        for p in range(self.world):
            ctx.buf[p] = hdl.buffer_ptrs[p]
            ctx.sig[p] = self.sig_hdl.buffer_ptrs[p]
            ctx.scratch[p] = self.scratch_hdl.buffer_ptrs[p]
        mc = 0
        try:
            mc = int(hdl.multicast_ptr) if hdl.has_multicast_support(self.device.type, self.device.index) else 0
        except Exception:
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        ctx.mc = mc
        ctx.rank, ctx.world = self.rank, self.world
        self.bufs[t.data_ptr()] = (t, hdl, ctx)
        return t

    def sub_ctx(self, t: torch.Tensor, offset: int) -> _CommCtx:
        """Context whose data pointers start ``offset`` elements into buffer ``t`` (a tail range)."""
        base = self._ctx(t)
        c = _CommCtx()
        for p in range(self.world):
            c.buf[p] = base.buf[p] + 4 * offset
            c.sig[p] = base.sig[p]
            c.scratch[p] = base.scratch[p]
        c.mc = (base.mc + 4 * offset) if base.mc else 0
        c.rank, c.world = self.rank, self.world
        return c

    def multicast_addr(self, t: torch.Tensor) -> int:
        return int(self._ctx(t).mc or 0)

    def barrier_on_stream(self, t: torch.Tensor):
        """Cross-rank barrier enqueued on the current stream (kernel-side flag barrier)."""
        _lib.call("fdt_comm_barrier", ctypes.byref(self._ctx(t)))

    def all_reduce_mean_ctx_(self, ctx: _CommCtx, n: int) -> torch.Tensor:
        _lib.call("fdt_comm_all_reduce_mean", ctypes.byref(ctx), _lib.i64(n), int(self.use_multicast))
        return self._take_sq()

    def owns(self, t: torch.Tensor) -> bool:
        return t.data_ptr() in self.bufs

    def _ctx(self, t: torch.Tensor) -> _CommCtx:
        return self.bufs[t.data_ptr()][2]

    def _take_sq(self) -> torch.Tensor:
        out = torch.empty((), dtype=torch.float32, device=self.device)
        _lib.call("fdt_comm_take_scalar", self.scratch, out)
        self.last_sqnorm = out
        return out

    # ------------------------------------------------------------------ collectives
    def all_reduce_mean_(self, flat: torch.Tensor) -> torch.Tensor:
        ctx = self._ctx(flat)
        _lib.call("fdt_comm_all_reduce_mean", ctypes.byref(ctx), _lib.i64(flat.numel()), int(self.use_multicast))
        self._take_sq()
        return flat

    def reduce_scatter_mean_(self, flat: torch.Tensor, lo: int, hi: int) -> torch.Tensor:
        ctx = self._ctx(flat)
        _lib.call("fdt_comm_reduce_scatter_mean", ctypes.byref(ctx), _lib.i64(lo), _lib.i64(hi), int(self.use_multicast))
        self._take_sq()
        return flat[lo:hi]

    def all_gather_(self, flat: torch.Tensor, lo: int, hi: int, shadow: Optional[torch.Tensor] = None) -> torch.Tensor:
        ctx = self._ctx(flat)
        _lib.call("fdt_comm_all_gather_cast", ctypes.byref(ctx), _lib.i64(hi - lo), shadow, int(self.use_multicast))
        return flat
