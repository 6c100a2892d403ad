"""CUDA execution path of :mod:`fdt.ops`: autograd wrappers around the sm_100a kernels in
``csrc/`` (loaded through :mod:`fdt.ops._lib`).

Tensor conventions on the GPU:
* CNN activations are bf16, logical ``[N, C, H, W]`` with channels_last strides, i.e. physically
  NHWC — what the implicit-GEMM conv kernels and the channel-wise kernels index as ``[P, C]``;
* conv / linear weights are fp32 master copies inside :class:`fdt.parallel.flat.FlatParameters`
  with a bf16 shadow refreshed once per step; conv weights are physically ``[O, kh, kw, I]``;
* the transformer's residual stream is fp32 ``[B, L, D]``; LayerNorm emits bf16 for the GEMMs.

Each fused unit saves exactly what its backward needs; the conv output ``y`` is kept (bf16)
rather than recomputed as the reference does (``resnet.py:108``) — with 180 GB of HBM the extra
conv per layer is the more expensive choice.
"""
from __future__ import annotations

import math
import os
from typing import Optional

import torch
import torch.nn.functional as F

from . import _lib
from ._lib import call, i64
from . import reference as ref

ACT = {"none": 0, "relu": 1, "celu": 2, "gelu": 3, "tanh": 4}
BF16 = torch.bfloat16

# Which implementation runs the GEMM-shaped work.  "tcgen05" = csrc/gemm_sm100.cu (hand-written
# tcgen05/TMEM/TMA kernels); "library" = cuDNN / cuBLAS through torch (kept as the oracle and for
# shapes the native kernels do not cover yet).
GEMM_BACKEND = os.environ.get("FDT_GEMM", "auto")


def _native_gemm() -> bool:
    if GEMM_BACKEND == "library":
        return False
    try:
        from . import gemm_sm100
        return gemm_sm100.available()
    except Exception:
        if GEMM_BACKEND == "tcgen05":
            raise
        return False


# ----------------------------------------------------------------------------- small helpers
def to_nhwc_bf16(x: torch.Tensor) -> torch.Tensor:
    return x.to(dtype=BF16, memory_format=torch.channels_last)


def _nhwc(x: torch.Tensor) -> torch.Tensor:
    if x.dtype != BF16 or not x.is_contiguous(memory_format=torch.channels_last):
        x = x.to(dtype=BF16, memory_format=torch.channels_last)
    return x


def _w16(p: torch.Tensor) -> torch.Tensor:
    """bf16 operand copy of a parameter: a view of the flat shadow when available."""
    f = getattr(p, "_fdt_flat", None)
    if f is not None:
        s = f[0].shadow_of(p)
        if s is not None:
            return s
    if p.dim() == 4:
        return p.detach().to(dtype=BF16, memory_format=torch.channels_last)
    return p.detach().to(BF16)


def _wdgrad(p: Optional[torch.Tensor], mode: int) -> Optional[torch.Tensor]:
    """Pre-transposed data-gradient operand of a flat-buffer parameter (FlatParameters.refresh_dgrad_shadow)."""
    f = getattr(p, "_fdt_flat", None) if p is not None else None
    return f[0].dgrad_shadow_of(p, mode) if f is not None else None


def cast_bf16_(src: torch.Tensor, dst: torch.Tensor):
    call("fdt_cast_bf16", src, dst, i64(src.numel()))


def sqnorm(g: torch.Tensor) -> torch.Tensor:
    out = torch.zeros((), dtype=torch.float32, device=g.device)
    call("fdt_sqnorm", g, i64(g.numel()), out)
    return out


def clip_scale_(g: torch.Tensor, sq: torch.Tensor, max_norm: float, skipped: Optional[torch.Tensor]):
    call("fdt_clip_scale", g, i64(g.numel()), sq, float(max_norm if max_norm is not None else 0.0), skipped)


def sgd_momentum_(p, d, buf, lr, momentum, dampening, nesterov, first):
    lr_ptr = lr if torch.is_tensor(lr) else None
    lr_val = 0.0 if torch.is_tensor(lr) else float(lr)
    call("fdt_sgd_momentum", p, d, buf, i64(p.numel()), lr_ptr, lr_val, float(momentum), float(dampening),
         int(bool(nesterov)), int(bool(first)))


def sgd_momentum_dev_(p, d, buf, hyper, nesterov, first, shadow=None):
    """hyper: device fp32 {lr, momentum, dampening} (graph-replay safe); ``shadow``: optional bf16 buffer that receives the
    updated parameters in the same pass."""
    call("fdt_sgd_momentum_dev", p, d, buf, i64(p.numel()), hyper, int(bool(nesterov)), int(bool(first)), shadow)


def clip_wd_(g, p, sq, max_norm, wd, skipped):
    call("fdt_clip_wd", g, p, i64(g.numel()), sq, float(max_norm if max_norm is not None else 0.0), float(wd), skipped)


def madgrad_dev_(p, g, v, s, x0, hyper, decay, eps, mirror, tick=True):
    call("fdt_madgrad_dev", p, g, v, s, x0, i64(p.numel()), hyper, float(decay), float(eps), int(bool(mirror)), int(bool(tick)),
         launches=2 if tick else 1)


def madgrad_(p, g, v, s, x0, lamb, ck, decay, eps, mirror):
    call("fdt_madgrad", p, g, v, s, x0, i64(p.numel()), float(lamb), float(ck), float(decay), float(eps), int(mirror))


_seed_state = {}
_SEED_ROWS = 1024


class _SeedTable:
    """Device-resident (seed, counter) pairs for the Philox streams of one device: a table of ``_SEED_ROWS`` rows with
    consecutive counters.  Every random kernel of a step takes the next row (a view — no clone, no launch); the Trainer
    advances the whole table by its length with ONE in-place add at the start of each step (also inside CUDA-graph
    replays), so no host synchronisation and no per-call bookkeeping kernels.  A backward pass regenerates its masks from
    the row its forward used; rows only change at step boundaries.  Without a Trainer the table is replaced (not mutated)
    when it runs out, so rows still held by pending backward passes stay valid."""

    def __init__(self, device):
        seed = torch.initial_seed() & 0x7FFFFFFFFFFF
        self.tab = torch.stack([torch.full((_SEED_ROWS,), seed, dtype=torch.int64), torch.arange(_SEED_ROWS, dtype=torch.int64)],
                               dim=1).to(device)
        self.idx = 0

    def next(self) -> torch.Tensor:
        if self.idx >= _SEED_ROWS:
            nxt = self.tab.clone()
            nxt[:, 1] += _SEED_ROWS
            self.tab, self.idx = nxt, 0
        row = self.tab[self.idx]
        self.idx += 1
        return row

    def begin_step(self):
        if self.idx:
            self.tab[:, 1] += _SEED_ROWS
            self.idx = 0


def reset_seed(device=None):
    """Restart the framework's Philox stream (dropout masks, augmentation, mixup draws) from ``torch.initial_seed()`` with
    counter 0 — call after ``torch.manual_seed`` to make two runs in one process draw the same sequence."""
    if device is None:
        _seed_state.clear()
    else:
        _seed_state.pop(torch.device(device), None)


def begin_step_seeds(device):
    """Advance the device's seed table (one launch) and hand out rows from its start again; called by the Trainer at the
    start of every step."""
    st = _seed_state.get(torch.device(device))
    if st is not None:
        st.begin_step()


def _next_seed_off(device) -> torch.Tensor:
    """Device-resident (seed, counter) pair for a Philox-driven kernel: the next row of the device's seed table."""
    device = torch.device(device)
    st = _seed_state.get(device)
    if st is None:
        st = _seed_state[device] = _SeedTable(device)
    return st.next()


def mixup_sample(alpha: Optional[float], B: Optional[int], device):
    """(lam, perm): lam ~ Beta(alpha, alpha) as a 0-d fp32 tensor and a random permutation of 0..B-1 (int64), drawn on the
    device by ONE tiny kernel from the framework's Philox stream; pass ``alpha=None`` / ``B=None`` to skip either."""
    so = _next_seed_off(device)
    lam = torch.empty((), dtype=torch.float32, device=device) if alpha is not None else None
    perm = torch.empty(int(B), dtype=torch.int64, device=device) if B else None
    call("fdt_mixup_sample", so, float(alpha if alpha is not None else 0.0), int(B or 0), lam, perm)
    return lam, perm


# ----------------------------------------------------------------------------- conv back ends
# BatchNorm batch statistics from the conv GEMM's epilogue (fp32 accumulators, before the bf16 rounding of y) instead
# of a separate pass over y.  FDT_BN_EPILOGUE_STATS=0 restores the separate reduction.
_EPI_STATS = os.environ.get("FDT_BN_EPILOGUE_STATS", "1") == "1"


class BNArena:
    """Per-model scratch for the batch statistics of every conv+BN unit: [Σy, Σy² | Σg, Σg·y] (4·C floats) per unit,
    cleared by ONE fill at the start of each forward pass instead of one memset per unit and direction.  ``gen`` lets
    a backward pass check that its slice was cleared for *its* forward (two forwards before one backward fall back
    to private scratch)."""

    def __init__(self):
        self.buf: Optional[torch.Tensor] = None
        self.gen = 0

    def begin(self, total: int, device) -> "BNArena":
        if self.buf is None or self.buf.numel() < total or self.buf.device != device:
            self.buf = torch.empty(total, dtype=torch.float32, device=device)
        self.buf.zero_()
        self.gen += 1
        return self

    def slice(self, off: int, n: int) -> torch.Tensor:
        return self.buf[off:off + n]


def _conv_fwd(x, w16, stride, padding, sums=None, keep=None):
    """``sums`` ([2*Cout] fp32, zero on entry): filled with Σy, Σy² per channel by the native paths.  ``keep``: list that
    receives the im2col matrix when the explicit path ran (reused by the backward pass).  Returns (y, sums_filled)."""
    if _native_gemm():
        from . import gemm_sm100
        y = gemm_sm100.conv_fwd(x, w16, stride, padding, sums)
        if y is None:       # stride 2 / stem / odd widths: explicit im2col + the same tcgen05 GEMM
            y = gemm_sm100.conv_fwd_im2col(x, w16, stride, padding, sums, keep)
        if y is not None:
            return y, sums is not None
    _lib.note_fallback("conv_fwd", f"cuDNN F.conv2d for x{tuple(x.shape)} w{tuple(w16.shape)} stride {stride}")
    return F.conv2d(x, w16, stride=stride, padding=padding), False


def _grad_slot(weight) -> Optional[torch.Tensor]:
    """The parameter's slice of the flat gradient buffer, as the physically contiguous
    [Cout, kh, kw, Cin] fp32 tensor the wgrad kernel can accumulate into directly."""
    if getattr(weight, "_fdt_flat", None) is None or weight.grad is None or weight.dim() != 4:
        return None
    g = weight.grad.permute(0, 2, 3, 1)
    return g if g.is_contiguous() and g.dtype == torch.float32 else None


def _flat_grad_slot(p) -> Optional[torch.Tensor]:
    """A 1-D / 2-D flat-buffer parameter's gradient view, for kernels that accumulate into it directly."""
    if p is None or getattr(p, "_fdt_flat", None) is None or p.grad is None or p.dim() > 2:
        return None
    g = p.grad
    return g if g.is_contiguous() and g.dtype == torch.float32 else None


def _mc_slot(weight):
    """(multicast address, alpha) when this parameter's gradient is all-reduced inside the wgrad
    kernel (DDP engine with an NVLS-mapped gradient buffer), else None."""
    return getattr(weight, "_fdt_mc_grad", None) if weight is not None else None


def _mc_accumulate(weight, dw_logical: torch.Tensor):
    """Fallback for a fused-reduce parameter whose gradient was produced by a library kernel: add it
    (scaled) into every rank's buffer with one multimem.red pass."""
    addr, alpha = weight._fdt_mc_grad
    src = dw_logical.float()
    if weight.dim() == 4:
        src = src.permute(0, 2, 3, 1)
    src = src.contiguous()
    call("fdt_comm_mc_red_add", _lib.ctypes.c_void_p(int(addr)), src, i64(src.numel()), float(alpha))


# Weight gradients that accumulate straight into the flat gradient buffer do not feed anything else in
# backward, and their split-K grids are small — so they run on a side stream, overlapping the data-gradient
# chain.  The operands are kept alive until `join_side_streams()` (called by the Trainer after backward,
# before the gradient reduction), which also makes the pattern safe inside CUDA-graph capture.
_WGRAD_SIDE = os.environ.get("FDT_WGRAD_STREAM", "0") == "1"      # opt-in: measured no gain on B200 (10.39 vs 10.37 ms/step)
_side = {"stream": None, "keep": [], "dirty": False}


def _wgrad_side_stream(device):
    if _side["stream"] is None or _side["stream"].device != device:
        _side["stream"] = torch.cuda.Stream(device)
    return _side["stream"]


def join_side_streams():
    if _side["dirty"]:
        ev = torch.cuda.Event()
        ev.record(_side["stream"])
        torch.cuda.current_stream().wait_event(ev)
        _side["dirty"] = False
    _side["keep"].clear()


def _conv_bwd(dy, x, w16, stride, padding, need_dx, need_dw, w_shape, weight=None, A=None, addend=None):
    """``addend``: gradient reaching x through the other branch of a residual fork; summed into dx in the dgrad epilogue."""
    dx = dw = None
    mc = _mc_slot(weight)
    if (_WGRAD_SIDE and need_dw and mc is None and _native_gemm() and weight is not None
            and _grad_slot(weight) is not None):
        from . import gemm_sm100, wgrad_sm100
        if wgrad_sm100._conv_ok(x.shape, w_shape, stride, padding):
            side = _wgrad_side_stream(dy.device)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            side.wait_event(ev)
            with torch.cuda.stream(side):
                gemm_sm100.conv_wgrad(dy, x, w_shape, stride, padding, _grad_slot(weight))
            _side["keep"].append((dy, x))
            _side["dirty"] = True
            need_dw = False
    if _native_gemm():
        from . import gemm_sm100
        if need_dx:
            dx = gemm_sm100.conv_dgrad(dy, w16, x.shape, stride, padding, _wdgrad(weight, 0), addend)
            if dx is not None:
                addend = None                   # already summed in the GEMM epilogue
            else:
                dx = gemm_sm100.conv_dgrad_im2col(dy, w16, x.shape, stride, padding, _wdgrad(weight, 1))
        if need_dw:
            if mc is not None:      # tile-by-tile in-switch all-reduce fused into the wgrad epilogue
                dw = gemm_sm100.conv_wgrad(dy, x, w_shape, stride, padding, mc[0], mc[1], True)
                if dw is None:
                    dw = gemm_sm100.conv_wgrad_im2col(dy, x, w_shape, stride, padding, mc[0], mc[1], True, A)
            else:
                slot = _grad_slot(weight) if weight is not None else None
                dw = gemm_sm100.conv_wgrad(dy, x, w_shape, stride, padding, slot)
                if dw is None:
                    dw = gemm_sm100.conv_wgrad_im2col(dy, x, w_shape, stride, padding, slot, A=A)
            if dw is True:          # accumulated straight into the flat gradient buffer(s)
                need_dw, dw = False, None
    if (need_dx and dx is None) or (need_dw and dw is None):
        mask = [need_dx and dx is None, need_dw and dw is None, False]
        _lib.note_fallback("conv_bwd", f"cuDNN convolution_backward (dx={mask[0]}, dw={mask[1]}) for x{tuple(x.shape)} w{tuple(w_shape)}")
        rdx, rdw, _ = torch.ops.aten.convolution_backward(dy, x, w16, None, [stride, stride], [padding, padding],
                                                          [1, 1], False, [0, 0], 1, mask)
        dx = rdx if dx is None else dx
        dw = rdw if dw is None else dw
    if dw is not None and mc is not None:
        _mc_accumulate(weight, dw)
        dw = None
    if addend is not None and dx is not None:
        _lib.note_fallback("conv_bwd", "residual-gradient sum as a separate ATen add (dgrad not on the implicit-GEMM kernel)")
        dx = dx + addend
    return dx, dw


# ----------------------------------------------------------------------------- conv + BN + act unit
# measured neutral on B200 (4.86 vs 4.83 ms forward+backward: these passes are not bound by the bytes they move), so off by default
_BWD_MASKY = os.environ.get("FDT_BN_BWD_MASKY", "0") == "1"


class _ConvBNAct(torch.autograd.Function):
    """y = conv(x, w);  out = act(bn(y) [+ residual]).  ``form`` 0: reference batch-stat form,
    1: BatchNorm2d training, 2: BatchNorm2d eval."""

    @staticmethod
    def forward(ctx, x, weight, gamma, beta, residual, cfg):
        stride, padding, eps, act, form = cfg["stride"], cfg["padding"], cfg["eps"], ACT[cfg["act"]], cfg["form"]
        x = _nhwc(x)
        w16 = _w16(weight)
        C = weight.shape[0]
        dev = x.device
        scratch = sums = None
        ctx.arena = None
        arena = cfg.get("arena")            # (BNArena, offset): statistics slots cleared once per forward pass
        if _EPI_STATS and form != 2 and C % 8 == 0:
            if arena is not None and arena[0].buf is not None and arena[0].buf.device == dev:
                slot = arena[0].slice(arena[1], 4 * C)
                sums = slot[:2 * C]
                ctx.arena = (arena[0], arena[0].gen, slot[2 * C:])
            else:
                scratch = torch.empty(5 * C, dtype=torch.float32, device=dev)      # sums[2C] | coef[3C]
                sums = scratch[:2 * C]
                sums.zero_()
        keep = [] if (torch.is_grad_enabled() and weight.requires_grad) else None
        y, presummed = _conv_fwd(x, w16, stride, padding, sums, keep)
        ctx.im2col_A = keep[0] if keep else None
        Nb, C, Ho, Wo = y.shape
        P = Nb * Ho * Wo
        bnctx = torch.empty(5 * C, dtype=torch.float32, device=dev)
        out = torch.empty_like(y)
        res = _nhwc(residual) if residual is not None else None
        if presummed:       # statistics came out of the conv epilogue: normalise + act with inline coefficients, one launch
            call("fdt_bn_act_fwd_fused", y, res, out, bnctx, sums, gamma, beta, cfg.get("running_mean"),
                 cfg.get("running_var"), i64(P), C, float(eps), float(cfg.get("momentum") or 0.0), form, act)
        else:
            if scratch is None:
                scratch = torch.empty(5 * C, dtype=torch.float32, device=dev)
            call("fdt_bn_act_fwd", y, res, out, bnctx, scratch, gamma, beta, cfg.get("running_mean"), cfg.get("running_var"),
                 i64(P), C, float(eps), float(cfg.get("momentum") or 0.0), form, act, launches=4 if form != 2 else 2)
        ctx.save_for_backward(x, weight, w16, y, out, bnctx)
        ctx.gamma, ctx.beta = gamma, beta
        ctx.cfg = (stride, padding, act, form, residual is not None, gamma is not None)
        if cfg.get("passthrough"):
            return out, x.view_as(x)            # alias of the input: its consumer's gradient comes back through this node
        return out

    @staticmethod
    def backward(ctx, gout, gpass=None):
        x, weight, w16, y, out, bnctx = ctx.saved_tensors
        stride, padding, act, form, has_res, affine = ctx.cfg
        Nb, C, Ho, Wo = y.shape
        P = Nb * Ho * Wo
        dev = y.device
        gout = _nhwc(gout)
        ar = getattr(ctx, "arena", None)
        if ar is not None and ar[0].gen == ar[1]:       # slot cleared by this step's forward pass: no memset
            sums, red, nl = ar[2], "fdt_chan_reduce2_acc", 1
        else:
            sums, red, nl = torch.empty(2 * C, dtype=torch.float32, device=dev), "fdt_chan_reduce2", 2
        # ReLU unit without a residual input: the mask is re-derived from y with the forward's coefficients in both passes,
        # so the activated output is not read and the masked gradient never stored (2 of the 7 tensor passes of the generic path)
        masky = (_BWD_MASKY and act == ACT["relu"] and not has_res and form in (0, 1) and C % 8 == 0 and 256 % (C // 8) == 0)
        if masky:
            g = gout
            call("fdt_chan_reduce2_masky", gout, y, bnctx, ctx.beta if affine else None, sums, int(nl == 1), i64(P), C, launches=nl)
        elif act != 0:
            g = torch.empty_like(gout)
            call(red, gout, y, out, g, sums, i64(P), C, act, launches=nl)
        else:
            g = gout
            call(red, gout, y, None, None, sums, i64(P), C, 0, launches=nl)
        dgamma = dbeta = None
        ret_gamma = ret_beta = None
        if affine:
            # written (not accumulated) by the kernel: straight into the parameters' slots of the flat gradient buffer when
            # there are any (a BatchNorm's gamma / beta are used by exactly one unit), else into fresh tensors
            dgamma, dbeta = _flat_grad_slot(ctx.gamma), _flat_grad_slot(ctx.beta)
            if dgamma is None or dbeta is None:
                dgamma = ret_gamma = torch.empty(C, dtype=torch.float32, device=dev)
                dbeta = ret_beta = torch.empty(C, dtype=torch.float32, device=dev)
        dy = torch.empty_like(y)
        if masky:
            call("fdt_bn_bwd_apply_masky", gout, y, dy, sums, bnctx, ctx.beta if affine else None, dgamma, dbeta, i64(P), C, form)
        else:
            call("fdt_bn_bwd_apply", g, y, dy, sums, bnctx, dgamma, dbeta, i64(P), C, form)
        addend = None
        if gpass is not None and ctx.needs_input_grad[0]:
            addend = _nhwc(gpass)
        dx, dw = _conv_bwd(dy, x, w16, stride, padding, ctx.needs_input_grad[0], ctx.needs_input_grad[1], weight.shape, weight,
                           getattr(ctx, "im2col_A", None), addend)
        return dx, dw, ret_gamma, ret_beta, (g if has_res else None), None


def conv_bnstat_act(x, weight, stride, padding, eps, act, residual, arena=None, passthrough=False):
    cfg = dict(stride=stride, padding=padding, eps=eps, act=act, form=0, arena=arena, passthrough=passthrough)
    return _ConvBNAct.apply(x, weight, None, None, residual, cfg)


def conv_bnaffine_act(x, weight, gamma, beta, running_mean, running_var, num_batches_tracked, stride, padding,
                      eps, momentum, training, act, residual, arena=None, passthrough=False):
    if training and num_batches_tracked is not None:
        num_batches_tracked.add_(1)
    cfg = dict(stride=stride, padding=padding, eps=eps, act=act, form=1 if training else 2,
               running_mean=running_mean, running_var=running_var, momentum=momentum, arena=arena, passthrough=passthrough)
    return _ConvBNAct.apply(x, weight, gamma, beta, residual, cfg)


class _PoolFC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        x = _nhwc(x)
        N, C, H, W = x.shape
        K = weight.shape[0]
        w = weight.detach().float().contiguous()
        pooled = torch.empty(N, C, dtype=torch.float32, device=x.device)
        logits = torch.empty(N, K, dtype=torch.float32, device=x.device)
        call("fdt_pool_fc_fwd", x, w, None if bias is None else bias.detach().float().contiguous(), pooled, logits,
             N, H * W, C, K)
        ctx.save_for_backward(w, pooled)
        ctx.weight, ctx.bias = weight, bias
        ctx.geom = (N, C, H, W, K, bias is not None)
        return logits

    @staticmethod
    def backward(ctx, g):
        w, pooled = ctx.saved_tensors
        N, C, H, W, K, has_bias = ctx.geom
        g = g.float().contiguous()
        dx = torch.empty((N, C, H, W), dtype=BF16, device=g.device, memory_format=torch.channels_last) \
            if ctx.needs_input_grad[0] else None
        # the kernel accumulates: straight into the flat gradient slots when the parameters have them
        wslot, bslot = _flat_grad_slot(ctx.weight), (_flat_grad_slot(ctx.bias) if has_bias else None)
        dW = wslot if wslot is not None else torch.zeros(K, C, dtype=torch.float32, device=g.device)
        db = None
        if has_bias:
            db = bslot if bslot is not None else torch.zeros(K, dtype=torch.float32, device=g.device)
        call("fdt_pool_fc_bwd", g, w, pooled, dx, dW, db, N, H * W, C, K)
        return dx, (None if wslot is not None else dW), (None if (bslot is not None or not has_bias) else db)


def pool_fc(x, weight, bias):
    """[N, C, H, W] bf16 NHWC → fp32 logits: global average pool + Linear in one kernel each way."""
    if weight.shape[0] > 32 or weight.shape[1] * 4 > 48 * 1024:
        _lib.note_fallback("pool_fc", f"ATen mean + cuBLAS linear for {tuple(weight.shape)} (native: <= 32 classes, <= 12288 channels)")
        return F.linear(x.float().mean(dim=(2, 3)), weight, bias)
    return _PoolFC.apply(x, weight, bias)


def input_aug_mixup(images_u8: torch.Tensor, train: bool, perm=None, lam=None, mean=None, std=None, pad: int = 4):
    """uint8 [B,H,W,3] → bf16 logical [B,3,H,W] (NHWC memory): crop + flip + normalise [+ mixup blend]."""
    B, H, W, C = images_u8.shape
    assert C == 3 and images_u8.dtype == torch.uint8
    out = torch.empty((B, 3, H, W), dtype=BF16, device=images_u8.device, memory_format=torch.channels_last)
    so = _next_seed_off(images_u8.device) if train else None
    lam_t = None
    if lam is not None:
        lam_t = lam.detach().reshape(-1).float().contiguous()
    import ctypes
    m3 = (ctypes.c_float * 3)(*mean)
    s3 = (ctypes.c_float * 3)(*std)
    call("fdt_input_aug_mixup", images_u8.contiguous(), perm, lam_t, 0 if lam_t is None else lam_t.numel(), so, out,
         B, H, W, pad, int(bool(train)), m3, s3)
    return out


class _InputAugMeta(torch.autograd.Function):
    """Meta-mixup on the raw uint8 batch (resnet50_test.py:388-401, 525): crop + flip + normalise both partners and blend
    them with the per-sample weight sigmoid(theta_b) in ONE kernel; the backward recomputes the partners from the uint8
    images (same Philox stream) and reduces dL/dx · (a − b) · sigmoid'(theta) to dtheta — the learnable-lambda path of
    ``--meta_learning`` with no torch ops and nothing saved but the seed."""

    @staticmethod
    def forward(ctx, images_u8, theta, perm, mean, std, pad):
        import ctypes
        B, H, W, C = images_u8.shape
        images_u8 = images_u8.contiguous()
        th = theta.detach().reshape(-1).float().contiguous()
        out = torch.empty((B, 3, H, W), dtype=BF16, device=images_u8.device, memory_format=torch.channels_last)
        so = _next_seed_off(images_u8.device)
        m3, s3 = (ctypes.c_float * 3)(*mean), (ctypes.c_float * 3)(*std)
        call("fdt_input_aug_mixup_ex", images_u8, perm, th, B, 1, so, out, B, H, W, pad, 1, m3, s3)
        ctx.save_for_backward(images_u8, th, perm, so)
        ctx.cfg = (tuple(mean), tuple(std), pad, theta.shape)
        return out

    @staticmethod
    def backward(ctx, g):
        import ctypes
        images_u8, th, perm, so = ctx.saved_tensors
        mean, std, pad, tshape = ctx.cfg
        B, H, W, _ = images_u8.shape
        g = _nhwc(g)
        dth = torch.zeros(B, dtype=torch.float32, device=g.device)
        m3, s3 = (ctypes.c_float * 3)(*mean), (ctypes.c_float * 3)(*std)
        call("fdt_input_aug_mixup_bwd", images_u8, perm, th, 1, so, g, dth, B, H, W, pad, 1, m3, s3)
        return None, dth.view(tshape), None, None, None, None


def input_aug_meta_mixup(images_u8, theta, perm, mean, std, pad: int = 4):
    """uint8 [B,H,W,3] + learnable theta [B,...] → blended bf16 batch, differentiable w.r.t. theta."""
    assert images_u8.dtype == torch.uint8 and images_u8.shape[-1] == 3
    return _InputAugMeta.apply(images_u8, theta, perm.long().contiguous(), tuple(mean), tuple(std), int(pad))


# ----------------------------------------------------------------------------- mixup
def mixup_blend(x, index, lam):
    lam = lam.to(x.dtype)
    if lam.dim() == 1:
        lam = lam.view(-1, *([1] * (x.dim() - 1)))
    return torch.lerp(x[index], x, lam)


class _MixupCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, y_a, y_b, lam):
        logits = logits.float().contiguous()
        B, C = logits.shape
        out = torch.zeros(2, dtype=torch.float32, device=logits.device)
        dlogits = torch.empty_like(logits)
        lshape = lam.shape
        lam_f = lam.detach().reshape(-1).float().contiguous()
        want_dlam = lam.requires_grad and lam_f.numel() == B
        dlam = torch.empty(B, dtype=torch.float32, device=logits.device) if want_dlam else None
        call("fdt_mixup_ce_ex", logits, y_a.contiguous(), y_b.contiguous(), lam_f, lam_f.numel(), out, dlogits, dlam, B, C)
        ctx.save_for_backward(dlogits, dlam)
        ctx.lshape = lshape
        return out[0]

    @staticmethod
    def backward(ctx, g):
        dlogits, dlam = ctx.saved_tensors
        return dlogits * g, None, None, (None if dlam is None else (dlam * g).view(ctx.lshape))


def mixup_cross_entropy(logits, y_a, y_b, lam):
    if not torch.is_tensor(lam):
        lam = torch.tensor([float(lam)], device=logits.device)
    if logits.shape[1] > 32 or (lam.requires_grad and lam.numel() != logits.shape[0]):
        _lib.note_fallback("mixup_cross_entropy", "ATen log_softmax path (> 32 classes, or a learnable scalar lambda)")
        return ref.mixup_cross_entropy(logits, y_a, y_b, lam)
    return _MixupCE.apply(logits, y_a, y_b, lam)


# ----------------------------------------------------------------------------- transformer ops
class _LayerNorm(torch.autograd.Function):
    """``passthrough``: also return the input as a second output.  In a pre-LN sub-layer x feeds both the LayerNorm and the
    residual add; when the residual consumes the alias, its gradient arrives here together with the LayerNorm's own and
    the kernel accumulates dx straight into it (the fp32 residual-stream gradient flows down the network in ONE buffer, no
    add kernels, no zero fills).  da / db accumulate directly into the parameters' slots of the flat gradient buffer."""

    @staticmethod
    def forward(ctx, x, a, b, eps, passthrough):
        shp = x.shape
        D = shp[-1]
        x2 = x.float().contiguous().view(-1, D)
        M = x2.shape[0]
        y = torch.empty(M, D, dtype=BF16, device=x.device)
        stats = torch.empty(M, 2, dtype=torch.float32, device=x.device)
        call("fdt_ln_fwd", x2, a, b, y, stats, M, D, float(eps))
        ctx.save_for_backward(x2, a, b, stats)
        ctx.eps, ctx.shp = eps, shp
        if passthrough:
            return y.view(shp), x.view_as(x)
        return y.view(shp)

    @staticmethod
    def backward(ctx, g, gpass=None):
        x2, a, b, stats = ctx.saved_tensors
        M, D = x2.shape
        g = g.to(BF16).contiguous().view(M, D)
        if gpass is not None and gpass.dtype == torch.float32 and gpass.is_contiguous():
            dx = gpass.view(M, D)                       # accumulate in place into the residual-stream gradient
        elif gpass is not None:
            dx = gpass.float().contiguous().view(M, D)
        else:
            dx = torch.zeros_like(x2)
        da_slot, db_slot = _flat_grad_slot(a), _flat_grad_slot(b)
        da = da_slot if da_slot is not None else torch.zeros(D, dtype=torch.float32, device=x2.device)
        db = db_slot if db_slot is not None else torch.zeros(D, dtype=torch.float32, device=x2.device)
        call("fdt_ln_bwd", x2, a, stats, g, dx, da, db, M, D, float(ctx.eps))
        return dx.view(ctx.shp), (None if da_slot is not None else da), (None if db_slot is not None else db), None, None


def layernorm_ref(x, a_2, b_2, eps, passthrough=False):
    if x.shape[-1] not in (128, 256, 512, 1024):
        _lib.note_fallback("layernorm", f"ATen ops for D={x.shape[-1]} (native: 128/256/512/1024)")
        y = ref.layernorm_ref(x, a_2, b_2, eps)
        return (y, x) if passthrough else y
    return _LayerNorm.apply(x, a_2, b_2, eps, bool(passthrough))


class _LinearLib(torch.autograd.Function):
    """bf16 GEMM with fp32 master weights: library back end (cuBLAS) with the same contract as the
    tcgen05 path (x any float dtype → bf16 out; fused bias + activation)."""

    @staticmethod
    def forward(ctx, x, weight, bias, act):
        x16 = x.to(BF16)
        w16 = _w16(weight)
        z = F.linear(x16, w16, None if bias is None else bias.to(BF16).reshape(-1))
        out = ref.activation(z, act) if act != "none" else z
        ctx.save_for_backward(x16, w16, z if act in ("gelu",) else out)
        ctx.act, ctx.has_bias, ctx.xdtype, ctx.bshape = act, bias is not None, x.dtype, None if bias is None else bias.shape
        return out

    @staticmethod
    def backward(ctx, g):
        x16, w16, saved = ctx.saved_tensors
        g = g.to(BF16)
        if ctx.act == "gelu":
            zf = saved.float()
            g = (g.float() * (0.5 * (1 + torch.erf(zf * 0.7071067811865476)) +
                              zf * 0.3989422804014327 * torch.exp(-0.5 * zf * zf))).to(BF16)
        elif ctx.act == "relu":
            g = g * (saved > 0)
        elif ctx.act == "tanh":
            g = (g.float() * (1 - saved.float() ** 2)).to(BF16)
        g2 = g.reshape(-1, g.shape[-1])
        dx = (g2 @ w16).view(x16.shape).to(ctx.xdtype) if ctx.needs_input_grad[0] else None
        dw = (g2.t() @ x16.reshape(-1, x16.shape[-1])).float() if ctx.needs_input_grad[1] else None
        db = g2.float().sum(0).view(ctx.bshape) if ctx.has_bias and ctx.needs_input_grad[2] else None
        return dx, dw, db, None


def linear(x, weight, bias, act, drop_p=0.0):
    """act(x Wᵀ + b), optionally followed by dropout(p) fused into the same elementwise pass (mask regenerated, not
    stored, in the backward)."""
    if _native_gemm():
        from . import gemm_sm100
        out = gemm_sm100.linear(x, weight, bias, act, drop_p)
        if out is not None:
            return out
    _lib.note_fallback("linear", f"cuBLAS for x[..., {x.shape[-1]}] -> {weight.shape[0]} (native: K, N multiples of 64)")
    out = _LinearLib.apply(x, weight, bias, act)
    return F.dropout(out, drop_p, True) if drop_p > 0 else out


def linear_dropout_add(x, weight, bias, residual, p):
    """residual + dropout(linear(x)); None when the fused path does not apply."""
    if not _native_gemm():
        return None
    from . import linear_sm100
    return linear_sm100.linear_dropout_add(x, weight, bias, residual, p)


def attention(q, k, v, key_mask, dropout_p, training, mask_compat):
    if _native_gemm():
        from . import gemm_sm100
        out = gemm_sm100.attention(q, k, v, key_mask, dropout_p if training else 0.0, mask_compat)
        if out is not None:
            return out
    _lib.note_fallback("attention", f"SDPA / ATen for q{tuple(q.shape)} (native: d_head 64, L <= 512)")
    if mask_compat:
        return ref.attention(q, k, v, key_mask, dropout_p, training, True)
    m = None if key_mask is None else (key_mask != 0)[:, None, None, :]
    return F.scaled_dot_product_attention(q, k, v, attn_mask=m, dropout_p=dropout_p if training else 0.0)


def qkv_attention(x, weights, biases, heads, key_mask, dropout_p, mask_compat):
    """One GEMM for Q | K | V (weights adjacent in the flat buffer) + attention reading the packed result in place."""
    if not _native_gemm():
        return None
    from . import attention_sm100, linear_sm100
    B, L, _ = x.shape
    if L > 512 or weights[0].shape[0] // heads != 64:
        return None
    qkv = linear_sm100.linear_fused(x, weights, biases)
    if qkv is None:
        return None
    return attention_sm100.attention_packed(qkv, heads, key_mask, dropout_p, mask_compat)


class _Embed(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tokens, types, tok_w, pos_w, seg_w, scale):
        B, L = tokens.shape
        D = tok_w.shape[1]
        tokens = tokens.long().contiguous()
        types = types.long().contiguous()
        out = torch.empty(B, L, D, dtype=torch.float32, device=tok_w.device)
        call("fdt_embed_fwd", tokens, types, tok_w, pos_w, seg_w, out, B * L, L, D, float(scale))
        ctx.save_for_backward(tokens, types, tok_w, pos_w, seg_w)
        ctx.shapes = (tok_w.shape, pos_w.shape, seg_w.shape, scale)
        return out

    @staticmethod
    def backward(ctx, g):
        tokens, types, tok_w, pos_w, seg_w = ctx.saved_tensors
        ts, ps, ss, scale = ctx.shapes
        B, L = tokens.shape
        D = ts[1]
        g = g.float().contiguous()
        # scatter-add straight into the tables' slots of the flat gradient buffer (cleared once per step): no 62 MB zero
        # fill of a dense [30522, 512] temporary and no dense accumulate pass for a gradient that touches <= B*L rows
        slots = [_flat_grad_slot(w) for w in (tok_w, pos_w, seg_w)]
        outs = [sl if sl is not None else torch.zeros(shp, dtype=torch.float32, device=g.device)
                for sl, shp in zip(slots, (ts, ps, ss))]
        call("fdt_embed_bwd", tokens, types, g, outs[0], outs[1], outs[2], B * L, L, D, float(scale))
        dtok, dpos, dseg = [None if sl is not None else o for sl, o in zip(slots, outs)]
        return None, None, dtok, dpos, dseg, None


def embed(tokens, token_types, tok_w, pos_w, seg_w, scale):
    return _Embed.apply(tokens, token_types, tok_w, pos_w, seg_w, scale)


def mlp_head(x, W1, b1, W2, b2, bias_grad_mean_compat):
    """Linear → ReLU (tcgen05 GEMM with fused bias+ReLU epilogue) → Linear(hidden → n_class): the narrow second layer runs on
    the small-N head kernel (``fdt_pool_fc_*`` with a 1-pixel "pool": logits, dX, dW, db in one launch each way)."""
    if bias_grad_mean_compat:
        B = x.shape[0]
        b1 = ref._ScaleGrad.apply(b1, 1.0 / B)
        b2 = ref._ScaleGrad.apply(b2, 1.0 / B)
    h = linear(x, W1, b1, "relu")
    n_out, hid = W2.shape
    if n_out <= 32 and hid * 4 <= 48 * 1024 and h.dtype == BF16:
        return _PoolFC.apply(h.reshape(-1, hid, 1, 1), W2, b2.reshape(-1))
    _lib.note_fallback("mlp_head", f"cuBLAS F.linear for the {n_out}-wide output layer (native: <= 32 outputs)")
    return F.linear(h.float(), W2, b2.reshape(-1))


class _ClsGather(torch.autograd.Function):
    """[B, L, D] fp32 residual stream → bf16 [B, D] first-token rows; the backward writes the whole [B, L, D] gradient
    (zero except row 0) in one pass — the buffer the residual-stream backward then accumulates into in place."""

    @staticmethod
    def forward(ctx, x):
        B, L, D = x.shape
        x = x.float().contiguous()
        out = torch.empty(B, D, dtype=BF16, device=x.device)
        call("fdt_cls_gather", x, out, B, L, D)
        ctx.geom = (B, L, D)
        return out

    @staticmethod
    def backward(ctx, g):
        B, L, D = ctx.geom
        dx = torch.empty(B, L, D, dtype=torch.float32, device=g.device)
        call("fdt_cls_scatter", g.float().contiguous(), dx, B, L, D)
        return dx


def cls_gather(x):
    if x.shape[-1] % 4:
        return x[:, 0, :].contiguous()
    return _ClsGather.apply(x)


class _HeadMix(torch.autograd.Function):
    """out = lam * drop(z) + (1 - lam) * drop(z[perm]): pooled-vector dropout + manifold mixup, one kernel each way."""

    @staticmethod
    def forward(ctx, z, perm, lam, p):
        B, D = z.shape
        z = z.to(BF16).contiguous()
        out = torch.empty_like(z)
        so = _next_seed_off(z.device) if p > 0 else None
        lam_t = None if lam is None else lam.detach().reshape(-1).float().contiguous()
        perm_t = None if perm is None else perm.long().contiguous()
        call("fdt_head_mix_fwd", z, perm_t, lam_t, out, B, D, float(p), so)
        ctx.save_for_backward(perm_t, lam_t, so)
        ctx.geom = (B, D, float(p), z.dtype)
        return out

    @staticmethod
    def backward(ctx, g):
        perm_t, lam_t, so = ctx.saved_tensors
        B, D, p, _ = ctx.geom
        dz = torch.empty(B, D, dtype=torch.float32, device=g.device)
        call("fdt_head_mix_bwd", g.to(BF16).contiguous(), perm_t, lam_t, dz, B, D, p, so, launches=2)
        return dz.to(BF16), None, None, None


def head_mix(z, perm, lam, p):
    if z.shape[-1] % 4 or (lam is not None and lam.numel() != 1):
        _lib.note_fallback("head_mix", "ATen dropout + lerp (per-sample lambda or odd width)")
        z = F.dropout(z, p, True) if p > 0 else z
        return z if perm is None else lam.to(z.dtype) * z + (1.0 - lam.to(z.dtype)) * z[perm]
    return _HeadMix.apply(z, perm, lam, float(p))


class _DropoutAdd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, p):
        x = x.contiguous()
        res = residual.float().contiguous()
        out = torch.empty_like(res)
        so = _next_seed_off(x.device)
        assert x.dtype in (BF16, torch.float32)
        call("fdt_dropout_add", x, int(x.dtype == BF16), res, out, i64(x.numel()), float(p), so)
        ctx.save_for_backward(so)
        ctx.p, ctx.xdtype = p, x.dtype
        return out

    @staticmethod
    def backward(ctx, g):
        (so,) = ctx.saved_tensors
        g = g.float().contiguous()
        dx = torch.empty(g.shape, dtype=ctx.xdtype, device=g.device)
        call("fdt_dropout_bwd", g, dx, int(ctx.xdtype == BF16), i64(g.numel()), float(ctx.p), so)
        return dx, g, None


def dropout_add(x, residual, p, training):
    if x.dtype not in (BF16, torch.float32):
        x = x.float()
    return _DropoutAdd.apply(x, residual, p if training else 0.0)


# ----------------------------------------------------------------------------- NGD
def ngd_eig_update(fb, K, L, trXXt, max_sweeps: int = 12):
    """(A, w_coeff, d1, rho1) for a FactorBatch — csrc/ngd_eig.cu."""
    if K.dtype != torch.float32:
        from ..optim.ngd import fisher_update_torch
        return fisher_update_torch(fb, None, K, L, trXXt)
    n, R = fb.n, fb.R
    dev = K.device
    A = torch.empty(n, R, R, dtype=torch.float32, device=dev)
    wc = torch.empty(n, R, dtype=torch.float32, device=dev)
    d1 = torch.empty(n, R, dtype=torch.float32, device=dev)
    rho1 = torch.empty(n, dtype=torch.float32, device=dev)
    call("fdt_ngd_eig_update", K.contiguous(), L.contiguous(), fb.d, fb.rho, trXXt.contiguous(), A, wc, d1, rho1,
         n, R, fb.D, float(fb.N), float(fb.alpha), float(fb.eta), int(max_sweeps))
    return A, wc, d1, rho1


_desc_cache = {}


def ngd_eig_update_multi(pending, max_sweeps: int = 12):
    """Refresh every factor of every pending FactorBatch with ONE launch (csrc/ngd_eig.cu).
    Inputs fb.K/L/d/rho/tr and outputs fb.A/wc/d1/rho1 are the batches' persistent buffers, so the
    96-byte per-factor descriptors are built once per set of batches and cached on the device."""
    import struct
    key = tuple(id(fb) for fb in pending)
    ent = _desc_cache.get(key)
    if ent is None:
        rows = []
        for fb in pending:
            R, D = fb.R, fb.D
            for f in range(fb.n):
                ptrs = (fb.K.data_ptr() + f * R * R * 4, fb.L.data_ptr() + f * R * R * 4, fb.d.data_ptr() + f * R * 4,
                        fb.rho.data_ptr() + f * 4, fb.tr.data_ptr() + f * 4, fb.A.data_ptr() + f * R * R * 4,
                        fb.wc.data_ptr() + f * R * 4, fb.d1.data_ptr() + f * R * 4, fb.rho1.data_ptr() + f * 4)
                rows.append((R, struct.pack("<9Q2i3fi", *ptrs, R, D, float(fb.N), float(fb.alpha), float(fb.eta), 0)))
        rows.sort(key=lambda r: -r[0])                     # biggest eigenproblems first
        blob = b"".join(r[1] for r in rows)
        host = torch.frombuffer(bytearray(blob), dtype=torch.uint8)
        dev = torch.empty(len(blob), dtype=torch.uint8, device=pending[0].K.device)
        dev.copy_(host)
        ent = (dev, len(rows), [fb for fb in pending])     # keep the batches (and their buffers) alive
        _desc_cache[key] = ent
    call("fdt_ngd_eig_update_multi", ent[0], ent[1], int(max_sweeps))


def ngd_norms_rescale(X: torch.Tensor, Xhat: torch.Tensor, sums: Optional[torch.Tensor] = None):
    """(p0, out): p0[f] = ΣX_f², out_f = X̂_f·sqrt(p0/(p1+1e-30)) (X_f if p1 is NaN), written in place into
    ``Xhat``.  Three launches (memset, reduce, rescale) instead of ~10 tensor ops."""
    n = X.shape[0]
    M = X[0].numel()
    if sums is None:            # else: the factor batch's slice of the optimizer's per-step arena (cleared by one fill)
        sums = torch.zeros(2 * n, dtype=torch.float32, device=X.device)
    call("fdt_ngd_sumsq2", X, Xhat, sums, n, i64(M))
    call("fdt_ngd_rescale", X, Xhat, sums, Xhat, n, i64(M))
    return sums[:n], Xhat


def ngd_small_axis_ok(D: int, R: int) -> bool:
    return 2 <= D <= 12 and R == (D + 1) // 2


class NormChain:
    """The factors of one tensor are applied axis after axis, and the reference restores the Frobenius norm after each of
    them (ngd_optimizer.py:168) — so every link's output has the norm of the chain's FIRST input, and since every link is
    linear in its input the intermediate rescales cancel: out = X̂_last · sqrt(Σ X_first² / Σ X̂_last²).  On steps without a
    Fisher refresh (whose statistics need the properly scaled input) the links therefore skip their rescale pass (a read
    and a write of the whole tensor each) and :meth:`finish` applies one at the end.  A non-finite link makes the final
    Σ non-finite and the chain returns its first input (the reference falls back per axis, ngd_optimizer.py:157-159)."""

    def __init__(self, x0: torch.Tensor):
        self.x0 = x0
        self.p_in = None
        self.p_out = None

    def link(self, sums: torch.Tensor, n: int):
        if self.p_in is None:
            self.p_in = sums[:n]
        self.p_out = sums[n:2 * n]

    def finish(self, xhat: torch.Tensor) -> torch.Tensor:
        if self.p_in is None:
            return xhat
        xh = xhat if xhat.is_contiguous() else xhat.contiguous()
        x0 = self.x0 if self.x0.is_contiguous() else self.x0.contiguous()
        n = xh.shape[0]
        call("fdt_ngd_rescale2", x0, xh, self.p_in, self.p_out, xh, n, i64(xh[0].numel()))
        return xh


def ngd_small_axis(X: torch.Tensor, W: torch.Tensor, A: int, D: int, B: int, J: Optional[torch.Tensor], chain=None):
    """Precondition a short axis of the physical gradient view ``X`` = [n, A, D, B] (csrc/ngd_axis.cu): projection,
    deflation, norm restoration and (when ``J`` is given) the Fisher statistics J = Σ h xᵀ, without materialising
    H or a transposed copy.  Returns (p0 [n], out [same shape as X]); with a :class:`NormChain` the rescale is left to
    the end of the chain."""
    n = X.shape[0]
    Xhat = torch.empty_like(X)
    sums = torch.empty(2 * n, dtype=torch.float32, device=X.device)
    call("fdt_ngd_small_axis", X, W, Xhat, sums, J, n, i64(A), int(D), i64(B), int(W.shape[1]), launches=1)
    if chain is not None:
        chain.link(sums, n)
    else:
        call("fdt_ngd_rescale", X, Xhat, sums, Xhat, n, i64(A * D * B))
    return sums[:n], Xhat


def ngd_small_matrix_ok(N: int, D: int, R: int) -> bool:
    return N * D + R * D + N * R <= 24576


def ngd_small_matrix(X: torch.Tensor, fb, left: bool, want_j: bool):
    """Whole-factor-in-shared-memory step (csrc/ngd_axis.cu::fdt_ngd_small_matrix).  X [n, N, D] or [n, D, N] (left);
    returns (p0 [n], preconditioned + rescaled X in the same layout); J goes to fb.Jp when ``want_j``."""
    n = X.shape[0]
    out = torch.empty_like(X)
    p0 = torch.empty(n, dtype=torch.float32, device=X.device)
    call("fdt_ngd_small_matrix", X, fb.Wp, fb.Dp, out, p0, fb.Jp if want_j else None, fb.Dp, n, fb.N, fb.D, fb.R, int(bool(left)))
    return p0, out


def ngd_kl_any(fb) -> bool:
    """K = J Jᵀ, L = J Wᵀ for a factor batch on whichever native kernel covers its shape; False if none does."""
    R, D = fb.R, fb.D
    if R % 4 == 0 and 16 <= R <= 128 and fb.Dp % 4 == 0:
        ngd_stats_kl(fb.Jp, fb.Wp, fb.K, fb.L)
        return True
    if R <= 16 and D <= 32 and fb.Dp == D:
        ngd_small_kl(fb.J, fb.W, fb.K, fb.L)
        return True
    return False


def ngd_small_kl(J, W, K, L):
    n, R, D = J.shape
    call("fdt_ngd_small_kl", J, W, K, L, n, R, D)


_commit_cache = {}


def ngd_commit_multi(pending):
    """Guarded hand-over W←W1, d←d1, rho←rho1 for every factor of every pending batch: one launch."""
    import struct
    key = tuple(id(fb) for fb in pending)
    ent = _commit_cache.get(key)
    if ent is None:
        rows = []
        for fb in pending:
            R, D = fb.R, fb.D
            for f in range(fb.n):
                D = fb.Dp
                rows.append(struct.pack("<6Q2i", fb.W1p.data_ptr() + f * R * D * 4, fb.d1.data_ptr() + f * R * 4,
                                        fb.rho1.data_ptr() + f * 4, fb.Wp.data_ptr() + f * R * D * 4,
                                        fb.d.data_ptr() + f * R * 4, fb.rho.data_ptr() + f * 4, R, D))
        blob = b"".join(rows)
        dev = torch.empty(len(blob), dtype=torch.uint8, device=pending[0].W.device)
        dev.copy_(torch.frombuffer(bytearray(blob), dtype=torch.uint8))
        flags = torch.zeros(len(rows), dtype=torch.int32, device=dev.device)
        ent = (dev, len(rows), list(pending), flags, max(fb.R * fb.Dp for fb in pending))
        _commit_cache[key] = ent
    call("fdt_ngd_commit_multi", ent[0], ent[1], ent[3], ent[4], launches=3)


_finish_cache = {}


def ngd_finish_multi(pending):
    """W1 = A (J + wc ∘ W) for every factor of every pending batch, then the guarded hand-over W←W1, d←d1,
    rho←rho1: three launches for the whole model (csrc/ngd_eig.cu::fdt_ngd_finish_multi)."""
    import struct
    key = tuple(id(fb) for fb in pending)
    ent = _finish_cache.get(key)
    if ent is None:
        rows = []
        for fb in pending:
            R, D = fb.R, fb.D
            for f in range(fb.n):
                Dp = fb.Dp            # the kernels walk the padded [R, Dp] storage (padding columns are and stay zero)
                rows.append(struct.pack(
                    "<9Q2i", fb.W1p.data_ptr() + f * R * Dp * 4, fb.d1.data_ptr() + f * R * 4, fb.rho1.data_ptr() + f * 4,
                    fb.Wp.data_ptr() + f * R * Dp * 4, fb.d.data_ptr() + f * R * 4, fb.rho.data_ptr() + f * 4,
                    fb.A.data_ptr() + f * R * R * 4, fb.Jp.data_ptr() + f * R * Dp * 4, fb.wc.data_ptr() + f * R * 4, R, Dp))
        blob = b"".join(rows)
        dev = torch.empty(len(blob), dtype=torch.uint8, device=pending[0].W.device)
        dev.copy_(torch.frombuffer(bytearray(blob), dtype=torch.uint8))
        flags = torch.zeros(len(rows), dtype=torch.int32, device=dev.device)
        ent = (dev, len(rows), list(pending), flags, max(fb.Dp for fb in pending), max(fb.R * fb.Dp for fb in pending))
        _finish_cache[key] = ent
    call("fdt_ngd_finish_multi", ent[0], ent[1], ent[3], ent[4], ent[5], launches=3)


# NGD's per-factor products on the tensor cores (csrc/ngd_gemm_sm100.cu): fp32-accurate 3xTF32 by default.
# FDT_NGD_TC=0 → batched cuBLAS fp32 (torch.bmm); FDT_NGD_TF32_TERMS=1 → plain tf32 (faster, 1e-3 relative).
NGD_TC = os.environ.get("FDT_NGD_TC", os.environ.get("FDT_NGD_TF32", "1")) == "1"
NGD_TF32 = NGD_TC       # old name


def ngd_tf32_ok(n, N, D, R, left) -> bool:
    """Shapes the batched tcgen05 NGD kernels cover; the rest stays on torch.bmm / the short-axis kernel.  With the
    gradient used transposed (``left``: the preconditioned axis is the tensor's first) D may be any length — W / J are
    stored with a padded row pitch; otherwise D is the contiguous axis of the gradient itself and must be a multiple of 4."""
    return NGD_TC and R % 4 == 0 and 16 <= R <= 128 and ((N % 4 == 0) if left else (D % 4 == 0))


def ngd_project(X: torch.Tensor, W: torch.Tensor, left: bool, sums: Optional[torch.Tensor] = None, D: Optional[int] = None,
                chain=None):
    """H = X Wᵀ and X̂ = X − H W for a factor batch on the tensor cores, then the norm-restoring rescale.  ``X`` is
    [n, N, D] (left=False) or the transposed gradient [n, D, N] (left=True), used where it lies.
    Returns (H [n, N, R], out (same layout as X), p0 [n]).  With a :class:`NormChain` the rescale is left to the end of
    the chain and ``out`` is the unscaled X̂."""
    n, R, Dw = W.shape              # W: padded storage [n, R, Dw] (Dw = D rounded up to 4, zero padding columns)
    D = Dw if D is None else D
    N = X[0].numel() // D
    H = torch.empty(n, N, R, dtype=torch.float32, device=X.device)
    Xhat = torch.empty_like(X)
    if sums is None:
        sums = torch.zeros(2 * n, dtype=torch.float32, device=X.device)
    call("fdt_ngd_project_ex", X, W, H, Xhat, sums, n, N, D, Dw, R, int(bool(left)), launches=2)
    if chain is not None:
        chain.link(sums, n)
    else:
        call("fdt_ngd_rescale", X, Xhat, sums, Xhat, n, i64(N * D))
    return H, Xhat, sums[:n]


def ngd_stats_j(H: torch.Tensor, X: torch.Tensor, J: torch.Tensor, left: bool, D: Optional[int] = None):
    """J = Hᵀ X on the tensor cores.  H [n, N, R]; X [n, N, D] (left=False) or [n, D, N] (left=True); J: padded storage
    [n, R, Dw]."""
    n, R, Dw = J.shape
    D = Dw if D is None else D
    N = H.shape[1]
    call("fdt_ngd_stats_j_ex", H, X, J, n, N, D, Dw, R, int(bool(left)))


def ngd_stats_hh(H: torch.Tensor, L: torch.Tensor):
    """L [n, R, R] = Hᵀ H (the N <= D case of the Fisher statistics)."""
    n, N, R = H.shape
    call("fdt_ngd_stats_hh", H, L, n, N, R)


def ngd_stats_kl(J: torch.Tensor, W: torch.Tensor, K: torch.Tensor, L: Optional[torch.Tensor]):
    """K = J Jᵀ and (if ``L`` is given) L = J Wᵀ on the tensor cores; J, W [n, R, D]; K, L [n, R, R]."""
    n, R, D = J.shape
    call("fdt_ngd_stats_kl", J, W, K, L, n, D, R, launches=2 if L is not None else 1)

