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
SIG_SLOTS = 2 * 128 * MAX_WORLD      # one channel: start + end barrier slots of up to 128 blocks x 8 ranks
SIG_CHANNELS = 4                     # 0: stand-alone collectives on the main stream; 1: bucketed collectives on the
                                     # communication stream (overlapping forward / backward); 2-3: free


class _CommCtx(ctypes.Structure):
    _fields_ = [("buf", ctypes.c_void_p * MAX_WORLD), ("sig", ctypes.c_void_p * MAX_WORLD),
                ("scratch", ctypes.c_void_p * MAX_WORLD), ("mc", ctypes.c_void_p),
                ("rank", ctypes.c_int), ("world", ctypes.c_int), ("sig_off", ctypes.c_int), ("pad1", ctypes.c_int)]


class SymmetricHeap:
    def __init__(self, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.symm_mem = symm_mem
        self.device = torch.device(device)
        self.group = group or dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        if self.world > MAX_WORLD:
            raise RuntimeError(f"symmetric heap supports up to {MAX_WORLD} ranks on one node")
        try:            # needed by older torch releases, a deprecated no-op in newer ones
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore", FutureWarning)
                symm_mem.enable_symm_mem_for_group(self.group.group_name)
        except Exception:
            pass
        self.sig = symm_mem.empty(SIG_SLOTS * SIG_CHANNELS, dtype=torch.int32, device=self.device)
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
        """A zero-filled fp32 buffer of ``numel`` elements that every rank can address on every peer.  The allocation is
        rounded up to a multiple of 4 x world elements (the kernels move 16-byte vectors and cut the buffer into one
        chunk per rank); the returned tensor is the first ``numel`` elements, and the collectives run over the padded
        length — the padding stays zero — so callers may use any length."""
        want = numel
        numel = (numel + 4 * self.world - 1) // (4 * self.world) * (4 * self.world)
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
        return t if want == numel else t[:want]

    def free(self, t: torch.Tensor):
        """Drop the heap's reference to a buffer from :meth:`alloc_f32` (the memory goes back once the caller's
        references are gone too).  Collective in spirit: every rank should free the same buffers."""
        self.bufs.pop(t.data_ptr(), None)
        for key in [k for k in getattr(self, "_chan", {}) if k[0] == t.data_ptr()]:
            self._chan.pop(key)

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

    def channel_ctx(self, t: torch.Tensor, channel: int) -> _CommCtx:
        """The buffer's context on another signal channel: collectives on different channels may be in flight at the
        same time (different streams) without touching each other's barrier flags."""
        assert 0 <= channel < SIG_CHANNELS
        key = (t.data_ptr(), channel)
        c = self._chan.get(key) if hasattr(self, "_chan") else None
        if c is None:
            if not hasattr(self, "_chan"):
                self._chan = {}
            base = self._ctx(t)
            c = _CommCtx()
            ctypes.memmove(ctypes.byref(c), ctypes.byref(base), ctypes.sizeof(_CommCtx))
            c.sig_off = channel * SIG_SLOTS
            self._chan[key] = c
        return c

    def reduce_scatter_range_(self, flat: torch.Tensor, lo: int, hi: int, blocks: int, channel: int = 1):
        """This rank's slice [lo, hi) of one gradient bucket (possibly empty); ``blocks`` is common to all ranks.
        Σg² keeps accumulating until :meth:`take_sq`."""
        _lib.call("fdt_comm_reduce_scatter_range", ctypes.byref(self.channel_ctx(flat, channel)), _lib.i64(lo), _lib.i64(hi),
                  int(blocks), int(self.use_multicast))

    def take_sq(self) -> torch.Tensor:
        return self._take_sq()

    def all_gather_range_(self, flat: torch.Tensor, sub_dev: torch.Tensor, shadow: Optional[torch.Tensor],
                          blocks: int, channel: int = 1):
        """Pull ``[sub_dev[2r], sub_dev[2r+1])`` (absolute element offsets, owned by rank r) from every owner, emitting
        the bf16 shadow of the same elements.  ``sub_dev``: int64 device tensor, identical on all ranks."""
        _lib.call("fdt_comm_all_gather_range", ctypes.byref(self.channel_ctx(flat, channel)), sub_dev, shadow, int(blocks))

    def bcast_range_(self, flat: torch.Tensor, owner: int, lo: int, hi: int, shadow: Optional[torch.Tensor], blocks: int,
                     channel: int = 1):
        """Two-phase broadcast of ``flat[lo:hi]`` from ``owner`` to every rank (scatter slices, then all-gather them), with the
        bf16 shadow of the same elements; for ranges so large that world-1 direct pulls would saturate the owner's links."""
        _lib.call("fdt_comm_bcast_range", ctypes.byref(self.channel_ctx(flat, channel)), int(owner), _lib.i64(lo), _lib.i64(hi),
                  shadow, int(blocks))

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
        n = self.bufs[flat.data_ptr()][0].numel()       # padded allocation length (>= flat.numel(), padding is zero)
        _lib.call("fdt_comm_all_reduce_mean", ctypes.byref(ctx), _lib.i64(n), int(self.use_multicast))
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
