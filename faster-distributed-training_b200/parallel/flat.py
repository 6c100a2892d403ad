"""Flat parameter / gradient storage.

All trainable tensors of a model are re-homed into ONE fp32 parameter buffer and ONE fp32
gradient buffer (``p.data`` / ``p.grad`` become views).  This is the B200-first replacement for
the reference's per-tensor world (69 / 105 separate tensors, each with its own all-reduce bucket
slot, ``clip_grad_norm_`` and optimizer launches — ``resnet50_test.py:499-502,546``,
``ngd_optimizer.py:462-506``):

* the DDP all-reduce, the FSDP reduce-scatter / all-gather, gradient clipping, zeroing and the
  bf16 weight-shadow cast each run over one contiguous range (one kernel, one collective);
* tensors with the **same shape are laid out adjacently**, so the NGD optimizer sees each shape
  group as a zero-copy ``[n, *shape]`` batch (see ``fdt/optim/ngd.py``);
* the gradient buffer can be placed in NVLink symmetric memory (``allocator``) so the fused
  all-reduce kernels in ``csrc/comm.cu`` can read peers' gradients directly;
* shape groups can be assigned to owner ranks (``owners``) and padded so every rank's region has
  the same length — the layout the FSDP engine shards on.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch

ALIGN = 128  # elements; every shape group (and rank region) starts on a 512-byte boundary


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


class ShapeGroupLayout:
    def __init__(self, shape, params, offset, owner=0):
        shape = tuple(d for d in shape if not isinstance(d, str))      # drop the "fused" bucket tag
        self.shape = tuple(shape)
        self.params = params
        self.numel = params[0].numel()
        self.offset = offset          # element offset of the first member in the flat buffers
        self.owner = owner            # FSDP owner rank (0 when unsharded)

    @property
    def n(self):
        return len(self.params)

    @property
    def size(self):
        return self.n * self.numel


class FlatParameters:
    def __init__(self, params: Sequence[torch.nn.Parameter], device=None,
                 allocator: Optional[Callable[[int], torch.Tensor]] = None,
                 world_size: int = 1, shard: bool = False,
                 param_allocator: Optional[Callable[[int], torch.Tensor]] = None,
                 fused_first: Optional[Callable[[torch.nn.Parameter], bool]] = None,
                 bucket_of: Optional[Callable[[torch.nn.Parameter], int]] = None,
                 equal_regions: bool = True):
        params = [p for p in params if p.requires_grad]
        assert params, "no trainable parameters"
        device = device or params[0].device
        self.device = torch.device(device)
        self.world_size = world_size
        self.sharded = shard and world_size > 1

        buckets: "OrderedDict[Tuple, List[torch.nn.Parameter]]" = OrderedDict()
        split_fused = fused_first is not None and not (shard and world_size > 1)
        for p in params:
            key = tuple(p.shape) + (("fused",) if split_fused and fused_first(p) else ())
            buckets.setdefault(key, []).append(p)
        if self.sharded:
            # a shape group larger than one rank's fair share is cut into parts of whole tensors, so that the regions
            # (all padded to the largest one) stay close to total / world and the optimizer work is balanced
            target = max(1, sum(p.numel() for p in params) // world_size)
            cut: "OrderedDict[Tuple, List[torch.nn.Parameter]]" = OrderedDict()
            for s, ps in buckets.items():
                per = max(1, target // max(1, ps[0].numel()))
                unit = int(getattr(ps[0], "_fdt_group_unit", 1))       # keep tensors that are used as one fused operand
                if unit > 1 and len(ps) % unit == 0:                   # (Q/K/V projection weights) in the same part
                    per = max(unit, per // unit * unit)
                if len(ps) <= per:
                    cut[s] = ps
                else:
                    for k in range(0, len(ps), per):
                        cut[s + (f"part{k // per}",)] = ps[k:k + per]
            buckets = cut
        # shape groups whose gradients are all-reduced inside the wgrad kernels go first, so that the
        # remaining ("tail") gradients form one contiguous range for the stand-alone all-reduce
        self.fused_end = 0
        n_fused = 0
        if fused_first is not None and not self.sharded:
            front = OrderedDict((s, ps) for s, ps in buckets.items() if s and s[-1] == "fused")
            n_fused = len(front)
            rest = OrderedDict((s, ps) for s, ps in buckets.items() if s not in front)
            buckets = OrderedDict(list(front.items()) + list(rest.items()))

        # owner assignment: greedy balance of element counts (largest groups first)
        owners: Dict[Tuple, int] = {s: 0 for s in buckets}
        if self.sharded:
            load = [0] * world_size
            for s, ps in sorted(buckets.items(), key=lambda kv: -len(kv[1]) * kv[1][0].numel()):
                r = min(range(world_size), key=lambda i: load[i])
                owners[s] = r
                load[r] += _round_up(len(ps) * ps[0].numel(), ALIGN)

        # Gradient buckets (sharded layout only): ``bucket_of(p)`` = index of the backward phase in which p's gradient
        # becomes final (0 = first to finish).  A shape group is ready with its LAST member; inside every rank's region the
        # groups are laid out bucket by bucket, so "bucket k" is one contiguous sub-range per rank —
        # ``bucket_ranges[k][r]`` — that a reduce-scatter can take as soon as phase k of the backward pass is over.
        gbucket: Dict[Tuple, int] = {s: (max(int(bucket_of(p)) for p in ps) if bucket_of is not None else 0)
                                     for s, ps in buckets.items()}
        self.num_buckets = (max(gbucket.values()) + 1) if gbucket else 1
        self.bucket_ranges: List[List[Tuple[int, int]]] = []
        self.groups: List[ShapeGroupLayout] = []
        self.region: List[Tuple[int, int]] = []   # per rank (start, end) element range
        off = 0
        if self.sharded:
            sizes = []
            for r in range(world_size):
                sz = sum(_round_up(len(ps) * ps[0].numel(), ALIGN) for s, ps in buckets.items() if owners[s] == r)
                sizes.append(sz)
            # equal_regions: every region padded to the largest one (what all_gather_into_tensor / reduce_scatter_tensor
            # and the single-launch native kernels need).  Otherwise regions are as long as what their rank owns — the
            # range collectives take per-rank bounds — so one very large tensor (a 30522 x 512 embedding) does not inflate
            # the buffers by world x its size.
            region_size = _round_up(max(sizes), ALIGN)
            self.bucket_ranges = [[(0, 0)] * world_size for _ in range(self.num_buckets)]
            starts = [r * region_size for r in range(world_size)]
            if not equal_regions:
                starts = [sum(sizes[:r]) for r in range(world_size)]
            for r in range(world_size):
                off = starts[r]
                self.region.append((off, off + (region_size if equal_regions else sizes[r])))
                for k in range(self.num_buckets):
                    lo = off
                    for s, ps in buckets.items():
                        if owners[s] != r or gbucket[s] != k:
                            continue
                        g = ShapeGroupLayout(s, ps, off, r)
                        g.bucket = k
                        self.groups.append(g)
                        off += _round_up(g.size, ALIGN)
                    self.bucket_ranges[k][r] = (lo, off)
            total = world_size * region_size if equal_regions else _round_up(sum(sizes), ALIGN)
        else:
            for gi, (s, ps) in enumerate(buckets.items()):
                g = ShapeGroupLayout(s, ps, off, 0)
                self.groups.append(g)
                off += _round_up(g.size, ALIGN)
                if gi + 1 == n_fused:
                    self.fused_end = off
            total = _round_up(off, ALIGN * max(world_size, 1))
            per = total // max(world_size, 1)
            self.region = [(r * per, (r + 1) * per) for r in range(max(world_size, 1))]
        self.total = total

        if param_allocator is not None:
            self.param_flat = param_allocator(total)
            self.param_flat.zero_()
        else:
            self.param_flat = torch.zeros(total, dtype=torch.float32, device=self.device)
        if allocator is not None:
            self.grad_flat = allocator(total)
            assert self.grad_flat.numel() == total and self.grad_flat.dtype == torch.float32
            self.grad_flat.zero_()
        else:
            self.grad_flat = torch.zeros(total, dtype=torch.float32, device=self.device)

        self._index: Dict[int, Tuple[ShapeGroupLayout, int]] = {}
        self.params: List[torch.nn.Parameter] = []
        self.shadow_flat: Optional[torch.Tensor] = None      # bf16 copy of param_flat, same layout
        with torch.no_grad():
            for g in self.groups:
                for i, p in enumerate(g.params):
                    lo = g.offset + i * g.numel
                    pv = self._view(self.param_flat, lo, g.shape)
                    pv.copy_(p.data.to(self.device, torch.float32))
                    p.data = pv
                    p.grad = self._view(self.grad_flat, lo, g.shape)
                    idx = len(self.params)
                    p._fdt_flat = (self, idx)
                    self._index[idx] = (g, i)
                    self.params.append(p)

    # 4-D (conv) weights are stored physically as [O, kh, kw, I] — the K-major operand layout of
    # the implicit-GEMM kernels and torch's channels_last — while their logical shape stays the
    # reference's [O, I, kh, kw], so state dicts are unchanged.
    @staticmethod
    def _view(flat: torch.Tensor, lo: int, shape, n: Optional[int] = None) -> torch.Tensor:
        numel = 1
        for s in shape:
            numel *= s
        lead = () if n is None else (n,)
        cnt = numel * (1 if n is None else n)
        if len(shape) == 4:
            O, I, H, W = shape
            v = flat[lo:lo + cnt].view(*lead, O, H, W, I)
            k = len(lead)
            return v.permute(*range(k), k, k + 3, k + 1, k + 2)
        return flat[lo:lo + cnt].view(*lead, *shape)

    def refresh_shadow(self):
        """bf16 shadow of every parameter in one pass (the GEMM / conv kernels' B operands)."""
        if self.device.type != "cuda":
            return
        if self.shadow_flat is None:
            self.shadow_flat = torch.empty(self.total, dtype=torch.bfloat16, device=self.device)
        if getattr(self, "shadow_fresh", False):
            self.shadow_fresh = False          # the optimizer's update kernel already wrote the shadow of what it changed
        else:
            from ..ops import cuda_ops
            cuda_ops.cast_bf16_(self.param_flat, self.shadow_flat)
        self.refresh_dgrad_shadow()

    # -- data-gradient operand layouts (transposed / tap-rotated weights), one launch per step
    def _build_dgrad_table(self):
        """Parameters flagged by their model with ``_fdt_dgrad_mode`` (0: implicit-GEMM dgrad / Linear,
        1: col2im dgrad) get a slot in one bf16 buffer that :meth:`refresh_dgrad_shadow` fills from the shadow."""
        import struct
        ents, total = [], 0
        fused_followers = set()
        for p in self.params:
            mates = getattr(p, "_fdt_fuse_with", None)
            if mates and p.dim() == 2 and self.adjacent_block([p] + list(mates)) is not None:
                fused_followers.update(id(m) for m in mates)
        for idx, p in enumerate(self.params):
            mode = getattr(p, "_fdt_dgrad_mode", None)
            if mode is None or p.dim() not in (2, 4) or id(p) in fused_followers:
                continue
            O, I = p.shape[0], p.shape[1]
            mates = getattr(p, "_fdt_fuse_with", None)
            if mates and all(id(m) in fused_followers for m in mates):
                O = O * (1 + len(mates))            # Q|K|V weights back to back = one [3*out, in] matrix → one [in, 3*out] transpose
            T = p.shape[2] * p.shape[3] if p.dim() == 4 else 1
            if O % 8 or I % 8:
                continue
            ents.append((idx, O, T, I, int(mode), total))
            total += O * T * I
        self._dgrad_views = {}
        self._dgrad_table = None
        if not ents:
            return
        self.dgrad_flat = torch.empty(total, dtype=torch.bfloat16, device=self.device)
        rows = []
        for idx, O, T, I, mode, off in ents:
            g, i = self._index[idx]
            src = self.shadow_flat.data_ptr() + 2 * (g.offset + i * g.numel)
            dst = self.dgrad_flat[off:off + O * T * I]
            rows.append(struct.pack("<2Q4i", src, dst.data_ptr(), O, T, I, mode))
            p = self.params[idx]
            if p.dim() == 2:
                view = dst.view(I, O)                                    # = W.t().contiguous()
            elif mode == 0:
                view = dst.view(I, p.shape[2], p.shape[3], O).permute(0, 3, 1, 2)   # [Cin, Cout, kh, kw] channels_last
            else:
                view = dst.view(T * I, O)                                # the [K, O] matrix of the col2im dgrad
            self._dgrad_views[idx] = (mode, view)
        blob = b"".join(rows)
        dev = torch.empty(len(blob), dtype=torch.uint8, device=self.device)
        dev.copy_(torch.frombuffer(bytearray(blob), dtype=torch.uint8))
        # one block per 64 x 64 tile: (descriptor, tile in that weight) for every block, so that no block searches the list
        per = [T * ((O + 63) // 64) * ((I + 63) // 64) for _, O, T, I, _, _ in ents]
        k = torch.repeat_interleave(torch.arange(len(per), dtype=torch.int32), torch.tensor(per))
        starts = torch.cumsum(torch.tensor([0] + per[:-1]), 0).to(torch.int32)
        local = torch.arange(sum(per), dtype=torch.int32) - starts[k.long()]
        tile_map = torch.stack([k, local], 1).contiguous().to(self.device)
        self._dgrad_table = (dev, tile_map, sum(per))

    def refresh_dgrad_shadow(self):
        if self.device.type != "cuda" or self.shadow_flat is None:
            return
        if not hasattr(self, "_dgrad_table"):
            self._build_dgrad_table()
        if self._dgrad_table is not None:
            from ..ops._lib import call
            call("fdt_dgrad_weights_map", self._dgrad_table[0], self._dgrad_table[1], self._dgrad_table[2])

    def dgrad_shadow_of(self, p: torch.nn.Parameter, mode: int) -> Optional[torch.Tensor]:
        ent = getattr(self, "_dgrad_views", {}).get(p._fdt_flat[1])
        if ent is None or ent[0] != mode or getattr(self, "_dgrad_table", None) is None:
            return None
        return ent[1]

    def shadow_of(self, p: torch.nn.Parameter) -> Optional[torch.Tensor]:
        if self.shadow_flat is None:
            return None
        g, i = self._index[p._fdt_flat[1]]
        return self._view(self.shadow_flat, g.offset + i * g.numel, g.shape)

    # -- zero-copy batched views for the optimizer
    def stacked_views(self, idxs: List[int], shape) -> Optional[Tuple[torch.Tensor, torch.Tensor]]:
        g0, i0 = self._index[idxs[0]]
        for k, idx in enumerate(idxs):
            g, i = self._index[idx]
            if g is not g0 or i != i0 + k:
                return None
        lo = g0.offset + i0 * g0.numel
        n = len(idxs)
        return (self._view(self.grad_flat, lo, shape, n), self._view(self.param_flat, lo, shape, n))

    def span_of(self, idxs: List[int]) -> Tuple[int, int]:
        """Element range [lo, hi) of the flat buffers covered by parameters ``idxs``."""
        lo = min(self._index[i][0].offset + self._index[i][1] * self._index[i][0].numel for i in idxs)
        hi = max(self._index[i][0].offset + (self._index[i][1] + 1) * self._index[i][0].numel for i in idxs)
        return lo, hi

    def adjacent_block(self, params) -> Optional[int]:
        """Element offset of ``params[0]`` when ``params`` (same shape) sit back to back in the flat buffers, i.e. form
        ONE [k * rows, ...] matrix in the parameter, gradient and bf16 shadow buffers alike; None otherwise."""
        try:
            ent = [self._index[p._fdt_flat[1]] for p in params]
        except (AttributeError, KeyError, TypeError):
            return None
        g0, i0 = ent[0]
        for k, (g, i) in enumerate(ent):
            if g is not g0 or i != i0 + k:
                return None
        return g0.offset + i0 * g0.numel

    @staticmethod
    def common_span(params) -> Optional[Tuple["FlatParameters", int, int]]:
        """(flat, lo, hi) when ``params`` all live in ONE FlatParameters buffer and no foreign parameter sits inside
        [lo, hi) — the range a flat optimizer kernel may sweep; None otherwise."""
        flats = {id(getattr(p, "_fdt_flat", (None,))[0]) for p in params}
        if len(flats) != 1 or not params or getattr(params[0], "_fdt_flat", None) is None:
            return None
        flat = params[0]._fdt_flat[0]
        idxs = [p._fdt_flat[1] for p in params]
        lo, hi = flat.span_of(idxs)
        mine = set(idxs)
        for i in range(len(flat.params)):
            if i not in mine and lo <= flat.span_of([i])[0] < hi:
                return None
        return flat, lo, hi

    def rebind_grads(self):
        """Re-point ``p.grad`` at the flat buffer (after someone set grads to ``None``)."""
        for idx, p in enumerate(self.params):
            g, i = self._index[idx]
            p.grad = self._view(self.grad_flat, g.offset + i * g.numel, g.shape)

    def zero_grad(self):
        self.grad_flat.zero_()

    def group_id_of(self, p: torch.nn.Parameter) -> int:
        return id(self._index[p._fdt_flat[1]][0])

    def owned_params(self, rank: int) -> List[torch.nn.Parameter]:
        return [p for g in self.groups if g.owner == rank for p in g.params]

    def local_region(self, rank: int) -> Tuple[int, int]:
        return self.region[rank]
