"""NGD — SGD whose gradient is preconditioned per tensor axis by an online natural-gradient
(low-rank-plus-identity inverse-Fisher) factor.

Capability parity with ``ngd_optimizer.py:8-508`` of the reference (``OnlineNaturalGradient`` +
``NGD``; maths in SURVEY.md §3.4), re-designed for a B200:

* **Batched by shape.**  Parameters with identical shapes are processed together: their
  gradients form one ``[n, *shape]`` tensor (a zero-copy view when the parameters live in a
  :class:`fdt.parallel.flat.FlatParameters` buffer) and every factor quantity is a batched
  tensor (``W [n,R,D]``, ``d [n,R]``, ``rho [n]``).  ResNet-50's 157 active factors become
  ~35 batched problems instead of 157 serial Python objects (``ngd_optimizer.py:474-491``).
* **No host synchronisation.**  The reference reads 7 scalars back per factor per update
  (``.item()`` at ``ngd_optimizer.py:225,240,285,287,289,300`` + ``math.isnan`` :158); here β, e,
  z-scale, floors, ρ and the NaN guard are device tensors, and on CUDA the R×R symmetric
  eigenproblem plus all the scalar algebra is ONE kernel launch for every factor of every
  group (``csrc/ngd_eig.cu``: in-shared-memory parallel Jacobi), so a step is capturable in a
  CUDA graph.
* **Tensor-only state** (``state_dict`` holds W/d/rho/momentum tensors and integer counters),
  unlike the reference whose state holds live Python objects.

Semantics kept exactly: weight decay is added to the gradient in place
*before* preconditioning; axes are preconditioned sequentially; the Frobenius norm of the
gradient is restored after each axis; at the first call the factor is default-initialised and
warm-started with 3 extra updates on the same gradient (``ngd_optimizer.py:356-376``); the Fisher
estimate is refreshed on every call for the first 10 calls, then every ``update_period``.
Fixed (Q18): Nesterov uses ``d_p + momentum * buf``.
"""
from __future__ import annotations

import math
import os
from collections import OrderedDict
from typing import Dict, List, Optional, Tuple

import torch
from torch.optim import Optimizer

EPSILON = 1.0e-10   # ngd_optimizer.py:37
DELTA = 5.0e-4      # ngd_optimizer.py:38
NUM_INITIAL_ITERS = 10


def default_rank(dim: int) -> int:
    return min((dim + 1) // 2, 80)


def orthonormal_special(rank: int, dim: int, dtype, device) -> torch.Tensor:
    """The deterministic near-orthonormal R×D start matrix (ngd_optimizer.py:397-420)."""
    first_elem = 1.1
    num_cols = dim // rank
    remainder = dim % rank
    k = torch.full((rank,), 1.0 / math.sqrt(first_elem * first_elem + num_cols - 1), dtype=dtype, device=device)
    k[:remainder] = 1.0 / math.sqrt(first_elem * first_elem + num_cols)
    diag = torch.diag(k)
    first = torch.diag(k * first_elem)
    return torch.cat((first, diag.repeat(1, num_cols + 1)), 1)[:, :dim].contiguous()


class FactorBatch:
    """State of ``n`` same-shaped Fisher factors acting on one axis of ``n`` same-shaped tensors."""

    def __init__(self, n: int, N: int, D: int, alpha: float, rank: int, update_period: int, eta: float,
                 dtype, device):
        assert D > 1
        if rank >= 0:
            assert 0 < rank < D
        else:
            rank = default_rank(D)
        self.n, self.N, self.D, self.R = n, N, D, rank
        self.alpha, self.update_period, self.eta = float(alpha), int(update_period), float(eta)
        self.t = 0
        E_tii = 1.0 / (2.0 + (D + rank) * self.alpha / D)
        W0 = math.sqrt(E_tii) * orthonormal_special(rank, D, dtype, device)
        # Row pitch Dp = D rounded up to 4 (zero padding columns): the tensor-core kernels need 16-byte-aligned rows, and
        # an axis such as the 30522-row token embedding is not a multiple of 4.  ``W`` is the [n, R, D] view everything
        # else uses; ``Wp`` the padded storage the native kernels address.
        self.Dp = D if D < 32 else (D + 3) // 4 * 4        # short axes never reach the tensor-core kernels
        self.Wp = torch.zeros(n, rank, self.Dp, dtype=dtype, device=device)
        self.W = self.Wp[:, :, :D]
        self.W.copy_(W0.unsqueeze(0))
        self.d = torch.full((n, rank), EPSILON, dtype=dtype, device=device)
        self.rho = torch.full((n,), EPSILON, dtype=dtype, device=device)

    # -- tensors that make up the checkpointable state
    def tensors(self) -> Dict[str, torch.Tensor]:
        return {"W": self.W, "d": self.d, "rho": self.rho}

    def updating(self) -> bool:
        return self.t < NUM_INITIAL_ITERS or self.t % self.update_period == 0


def fisher_update_torch(fb: FactorBatch, J, K, L, trXXt):
    """One refresh of (W, d, rho) from the statistics J=HᵀX, K=JJᵀ, L — pure torch, batched,
    sync-free.  Mirrors ngd_optimizer.py:225-324; also the oracle for ``csrc/ngd_eig.cu``."""
    n, N, D, R = fb.n, fb.N, fb.D, fb.R
    alpha, eta = fb.alpha, fb.eta
    W, d, rho = fb.W, fb.d, fb.rho
    d_sum = d.sum(1)
    beta = rho * (1.0 + alpha) + alpha * d_sum / D
    e = 1.0 / (beta[:, None] / d + 1.0)
    inv_sqrt_e = 1.0 / torch.sqrt(e)
    z_scale = torch.diagonal(K, dim1=1, dim2=2).sum(1).clamp_min(1.0)
    dpr = d + rho[:, None]
    eN = eta / N
    c1 = (eN * eN) / z_scale
    c2 = (eN * (1.0 - eta)) / z_scale
    c3 = ((1.0 - eta) ** 2) / z_scale
    outer0 = inv_sqrt_e[:, :, None] * inv_sqrt_e[:, None, :]
    op1 = inv_sqrt_e[:, :, None] * (inv_sqrt_e * dpr)[:, None, :]
    Z = K * outer0 * c1[:, None, None] + L * (op1 + op1.transpose(1, 2)) * c2[:, None, None]
    Z = Z + torch.diag_embed(c3[:, None] * dpr * dpr)
    Z = 0.5 * (Z + Z.transpose(1, 2))
    c, U = torch.linalg.eigh(Z)
    c = c.flip(1)
    U = U.flip(2)
    # the reference builds this floor as a default-dtype (fp32) tensor (ngd_optimizer.py:278);
    # rounding it the same way keeps fp64 trajectories bit-comparable and is a no-op in fp32
    c_floor = ((rho * (1.0 - eta)) ** 2 / z_scale).float().to(c.dtype)
    c = torch.maximum(c, c_floor[:, None])
    sqrt_c = c.sqrt() * z_scale.sqrt()[:, None]
    inv_sqrt_c = 1.0 / sqrt_c
    rho1 = (1.0 / (D - R)) * (eN * trXXt + (1.0 - eta) * (D * rho + d_sum) - sqrt_c.sum(1))
    floor_val = torch.clamp_min(DELTA * sqrt_c.max(1).values, EPSILON)
    d1 = torch.maximum(sqrt_c - rho1[:, None], floor_val[:, None])
    rho1 = torch.maximum(rho1, floor_val)
    beta1 = rho1 * (1.0 + alpha) + alpha * d1.sum(1) / D
    e1 = 1.0 / (beta1[:, None] / d1 + 1.0)
    sqrt_e1 = torch.sqrt(e1)
    w_coeff = ((1.0 - eta) / eN) * dpr
    left = eN * sqrt_e1 * inv_sqrt_c
    A = U.transpose(1, 2) * (left[:, :, None] * inv_sqrt_e[:, None, :])
    return A, w_coeff, d1, rho1


def _fisher_update(fb: FactorBatch, J, K, L, trXXt):
    if K.is_cuda:
        from .. import ops
        if ops.use_native(K):
            return ops._cu().ngd_eig_update(fb, K, L, trXXt)
    return fisher_update_torch(fb, J, K, L, trXXt)


def _split_factor(N: int, target: int = 2048) -> int:
    """Split count for the long reduction J = HᵀX (over the N rows).  Skinny factors (e.g. the 3-wide
    kernel axes of a conv weight: N ~ 8e5, D = 3, R = 2) would otherwise run as a single CTA walking
    the whole reduction; S independent partial products summed afterwards keep the GPU busy."""
    if N <= 4 * target:
        return 1
    S = N // target
    while S > 1 and N % S:
        S -= 1
    return S


def _alloc_stats(fb: FactorBatch):
    """Persistent work buffers of a factor batch's Fisher refresh (stable addresses for graph replay)."""
    n, R, D = fb.n, fb.R, fb.D
    kw = dict(dtype=fb.W.dtype, device=fb.W.device)
    fb.Jp = torch.zeros(n, R, fb.Dp, **kw)
    fb.J = fb.Jp[:, :, :D]
    fb.K = torch.empty(n, R, R, **kw)
    fb.L = torch.empty(n, R, R, **kw)
    fb.A = torch.empty(n, R, R, **kw)
    fb.W1p = torch.zeros(n, R, fb.Dp, **kw)
    fb.W1 = fb.W1p[:, :, :D]
    fb.wc = torch.empty(n, R, **kw)
    fb.d1 = torch.empty(n, R, **kw)
    fb.rho1 = torch.empty(n, **kw)
    fb.tr = torch.empty(n, **kw)


def _statistics(fb: FactorBatch, Xm: torch.Tensor, left: bool, H: torch.Tensor):
    """J = HᵀX, L, K = JJᵀ into the factor batch's persistent work buffers."""
    W = fb.W
    if not hasattr(fb, "J"):
        _alloc_stats(fb)
    if getattr(fb, "_h_rows", False):
        # H is [n, N, R] for both layouts (tcgen05 path): J on the tensor cores, L / K are R x R
        from .. import ops
        cu = ops._cu()
        cu.ngd_stats_j(H, Xm, fb.Jp, left, fb.D)
        if fb.N > fb.D:
            cu.ngd_stats_kl(fb.Jp, fb.Wp, fb.K, fb.L)     # K = J Jᵀ, L = J Wᵀ
        else:
            cu.ngd_stats_hh(H, fb.L)                      # L = Hᵀ H
            cu.ngd_stats_kl(fb.Jp, fb.Wp, fb.K, None)
        return
    S = _split_factor(fb.N)
    if left:   # Xm: [n, D, N], H: [n, R, N];  J = H Xᵀ reduces over N
        if S > 1:
            n, R, D, c = fb.n, fb.R, fb.D, fb.N // S
            Hs = H.reshape(n, R, S, c).transpose(1, 2).reshape(n * S, R, c)
            Xs = Xm.reshape(n, D, S, c).transpose(1, 2).reshape(n * S, D, c)
            torch.sum(torch.bmm(Hs, Xs.transpose(1, 2)).view(n, S, R, D), dim=1, out=fb.J)
        else:
            torch.bmm(H, Xm.transpose(1, 2), out=fb.J)
        if fb.N > fb.D:
            torch.bmm(fb.J, W.transpose(1, 2), out=fb.L)
        else:
            torch.bmm(H, H.transpose(1, 2), out=fb.L)
    else:      # Xm: [n, N, D], H: [n, N, R];  J = Hᵀ X reduces over N
        if S > 1:
            n, R, D, c = fb.n, fb.R, fb.D, fb.N // S
            Hs = H.reshape(n * S, c, R)
            Xs = Xm.reshape(n * S, c, D)
            torch.sum(torch.bmm(Hs.transpose(1, 2), Xs).view(n, S, R, D), dim=1, out=fb.J)
        else:
            torch.bmm(H.transpose(1, 2), Xm, out=fb.J)
        if fb.N > fb.D:
            torch.bmm(fb.J, W.transpose(1, 2), out=fb.L)
        else:
            torch.bmm(H.transpose(1, 2), H, out=fb.L)
    torch.bmm(fb.J, fb.J.transpose(1, 2), out=fb.K)


def _apply_refresh(fb: FactorBatch, A, w_coeff, d1, rho1):
    W = fb.W
    B = torch.addcmul(fb.J, w_coeff.unsqueeze(2), W)
    W1 = torch.bmm(A, B)
    # device-side guard: a factor whose refresh produced non-finite numbers keeps its old state
    ok = torch.isfinite(W1).flatten(1).all(1) & torch.isfinite(d1).all(1) & torch.isfinite(rho1)
    # in-place so that a captured CUDA graph keeps reading / writing the same state buffers
    fb.W.copy_(torch.where(ok[:, None, None], W1, W))
    fb.d.copy_(torch.where(ok[:, None], d1, fb.d))
    fb.rho.copy_(torch.where(ok, rho1, fb.rho))


def _refresh(fb: FactorBatch, Xm: torch.Tensor, left: bool, H: torch.Tensor, trXXt: torch.Tensor):
    """Update (W, d, rho) in place from the current gradient matrix and H (= X Wᵀ or its
    transpose)."""
    _statistics(fb, Xm, left, H)
    A, w_coeff, d1, rho1 = _fisher_update(fb, fb.J, fb.K, fb.L, trXXt)
    _apply_refresh(fb, A, w_coeff, d1, rho1)


def flush_refreshes(pending: List[FactorBatch]):
    """Finish the deferred Fisher refreshes of a whole optimizer step.  The statistics (J, K, L,
    tr) of every pending factor batch are already in its work buffers; on CUDA all eigenproblems —
    every factor of every shape group — run in ONE kernel launch (one CTA per factor)."""
    if not pending:
        return
    fb0 = pending[0]
    native = False
    if fb0.K.is_cuda and fb0.K.dtype == torch.float32:
        from .. import ops
        native = ops.use_native(fb0.K)
    if native:
        from .. import ops
        cu = ops._cu()
        cu.ngd_eig_update_multi(pending)
        cu.ngd_finish_multi(pending)      # W1 = A (J + w_coeff ∘ W), then the guarded W←W1, d←d1, rho←rho1
    else:
        for fb in pending:
            A, w_coeff, d1, rho1 = fisher_update_torch(fb, fb.J, fb.K, fb.L, fb.tr)
            _apply_refresh(fb, A, w_coeff, d1, rho1)
    pending.clear()


def precondition(fb: FactorBatch, Xm: torch.Tensor, left: bool, pending: Optional[List[FactorBatch]] = None,
                 chain=None) -> torch.Tensor:
    """Apply (and, on refresh steps, update) the inverse-Fisher factor.

    ``Xm`` is ``[n, N, D]`` (``left=False``) or its transpose ``[n, D, N]`` (``left=True``, used
    when the preconditioned axis is the tensor's first axis so no transpose copy is needed).
    Returns the preconditioned matrix in the same layout with its Frobenius norm restored
    (ngd_optimizer.py:138-168).

    The output only depends on the *current* W (the reference also projects with W_t before it
    computes W_{t+1}, ngd_optimizer.py:195-203), so when ``pending`` is given the refresh is
    deferred: its statistics are computed here and the factor batch is appended to ``pending``
    for :func:`flush_refreshes`."""
    native = False
    if Xm.is_cuda and Xm.dtype == torch.float32:
        from .. import ops
        native = ops.use_native(Xm)
    Xc = Xm.contiguous() if native else Xm
    p0 = None if native else (Xm * Xm).sum((1, 2))

    def project():
        W = fb.W
        if left:
            H = torch.bmm(W, Xm)
            return H, torch.baddbmm(Xm, W.transpose(1, 2), H, alpha=-1.0)
        H = torch.bmm(Xm, W.transpose(1, 2))
        return H, torch.baddbmm(Xm, H, W, alpha=-1.0)

    tc = False
    from .. import ops
    if native:
        tc = ops._cu().ngd_tf32_ok(fb.n, fb.N, fb.D, fb.R, left)
    if fb.t == 0:
        # warm start: 3 refreshes on the same gradient (ngd_optimizer.py:356-376)
        if tc:          # same tensor-core kernels as the steady state: no library GEMM anywhere in the optimizer
            from .. import ops
            fb._h_rows = True
            for _ in range(3):
                H, _, p0w = ops._cu().ngd_project(Xc, fb.Wp, left, None, fb.D)
                _refresh(fb, Xc, left, H, p0w)
        elif native and fb.W.dtype == torch.float32 and ops._cu().ngd_small_matrix_ok(fb.N, fb.D, fb.R):
            cu = ops._cu()
            _alloc_stats(fb)
            for _ in range(3):
                p0s, _ = cu.ngd_small_matrix(Xc, fb, left, True)
                if not cu.ngd_kl_any(fb):
                    torch.bmm(fb.J, fb.W.transpose(1, 2), out=fb.L)
                    torch.bmm(fb.J, fb.J.transpose(1, 2), out=fb.K)
                fb.tr.copy_(p0s)
                A_, w_coeff, d1, rho1 = _fisher_update(fb, fb.J, fb.K, fb.L, fb.tr)
                _apply_refresh(fb, A_, w_coeff, d1, rho1)
        else:
            if p0 is None:
                p0 = (Xm * Xm).sum((1, 2))
            fb._h_rows = False
            for _ in range(3):
                H, _ = project()
                _refresh(fb, Xm, left, H, p0)
    updating = fb.updating()
    fb.t += 1
    if native:
        from .. import ops
        cu = ops._cu()
        if tc:
            fb._h_rows = True
            # tcgen05 3xTF32: H, X̂ (+Σx², Σx̂²), rescale (left to the end of the tensor's chain of axes when there is no refresh)
            H, out, p0 = cu.ngd_project(Xc, fb.Wp, left, getattr(fb, "_sums", None), fb.D, None if updating else chain)
            if updating:
                if pending is None:
                    _refresh(fb, Xc, left, H, p0)
                else:
                    _statistics(fb, Xc, left, H)
                    fb.tr.copy_(p0)
                    pending.append(fb)
            return out
    if native and fb.W.dtype == torch.float32:
        from .. import ops
        cu = ops._cu()
        if cu.ngd_small_matrix_ok(fb.N, fb.D, fb.R):
            if updating and not hasattr(fb, "J"):
                _alloc_stats(fb)
            p0s, out = cu.ngd_small_matrix(Xc, fb, left, updating)
            if updating:
                if not cu.ngd_kl_any(fb):             # L = J Wᵀ (= Hᵀ H exactly, since H = X Wᵀ)
                    torch.bmm(fb.J, fb.W.transpose(1, 2), out=fb.L)
                    torch.bmm(fb.J, fb.J.transpose(1, 2), out=fb.K)
                fb.tr.copy_(p0s)
                if pending is None:
                    A_, w_coeff, d1, rho1 = _fisher_update(fb, fb.J, fb.K, fb.L, fb.tr)
                    _apply_refresh(fb, A_, w_coeff, d1, rho1)
                else:
                    pending.append(fb)
            return out
    H, Xhat = project()
    if native:
        from .. import ops
        from ..ops import _lib
        _lib.note_fallback("ngd_precondition", f"cuBLAS fp32 bmm for a factor with N={fb.N}, D={fb.D}, R={fb.R} (native: R % 4 == 0, "
                                               f"16 <= R <= 128, contiguous axis % 4 == 0, or D <= 12)")
        p0, out = ops._cu().ngd_norms_rescale(Xc, Xhat, getattr(fb, "_sums", None))      # fused Σx², Σx̂², rescale + NaN guard
        if updating:
            if pending is None:
                _refresh(fb, Xm, left, H, p0)
            else:
                _statistics(fb, Xm, left, H)
                fb.tr.copy_(p0)
                pending.append(fb)
        return out
    if updating:
        if pending is None:
            _refresh(fb, Xm, left, H, p0)
        else:
            _statistics(fb, Xm, left, H)
            fb.tr.copy_(p0)
            pending.append(fb)
    p1 = (Xhat * Xhat).sum((1, 2))
    scale = torch.sqrt(p0 / (p1 + 1.0e-30))
    out = Xhat * scale[:, None, None]
    return torch.where(torch.isnan(p1)[:, None, None], Xm, out)


def _small_axis_native(fb: FactorBatch, d_p: torch.Tensor) -> bool:
    if not (d_p.is_cuda and d_p.dtype == torch.float32):
        return False
    from .. import ops
    return ops.use_native(d_p) and ops._cu().ngd_small_axis_ok(fb.D, fb.R)


def precondition_small_axis(fb: FactorBatch, d_p: torch.Tensor, pa: int, pending: Optional[List[FactorBatch]],
                            chain=None) -> torch.Tensor:
    """:func:`precondition` for a short axis (D <= 12) of the stacked physical tensor ``d_p`` = [n, *phys]: the axis
    at physical position ``pa`` is preconditioned in place of the ``[n, N, D]`` matrix view, so no transposed copy
    is made and H is never materialised (csrc/ngd_axis.cu).  Same math as ngd_optimizer.py:138-168,195-218; the
    statistics use L = J Wᵀ (N > D always holds here, ngd_optimizer.py:205-210)."""
    from .. import ops
    cu = ops._cu()
    n = d_p.shape[0]
    phys = d_p.shape[1:]
    A = int(math.prod(phys[:pa]))
    B = int(math.prod(phys[pa + 1:]))
    X = d_p.contiguous()
    if fb.t == 0:
        # warm start (ngd_optimizer.py:356-376): 3 refreshes on the same gradient, on the same native kernels
        _alloc_stats(fb)
        for _ in range(3):
            p0w, _ = cu.ngd_small_axis(X, fb.W, A, fb.D, B, fb.J)
            cu.ngd_small_kl(fb.J, fb.W, fb.K, fb.L)
            fb.tr.copy_(p0w)
            A_, w_coeff, d1, rho1 = _fisher_update(fb, fb.J, fb.K, fb.L, fb.tr)
            _apply_refresh(fb, A_, w_coeff, d1, rho1)
    updating = fb.updating()
    fb.t += 1
    # (N <= D: the reference takes L = Hᵀ H there, ngd_optimizer.py:215-217 — the same matrix as J Wᵀ, since H = X Wᵀ)
    if updating and not hasattr(fb, "J"):
        _alloc_stats(fb)
    p0, out = cu.ngd_small_axis(X, fb.W, A, fb.D, B, fb.J if updating else None, None if updating else chain)
    if updating:
        cu.ngd_small_kl(fb.J, fb.W, fb.K, fb.L)      # K = J Jᵀ, L = J Wᵀ in one small launch
        fb.tr.copy_(p0)
        if pending is None:
            A_, w_coeff, d1, rho1 = _fisher_update(fb, fb.J, fb.K, fb.L, fb.tr)
            _apply_refresh(fb, A_, w_coeff, d1, rho1)
        else:
            pending.append(fb)
    return out


class _ShapeGroup:
    """Same-shaped parameters of one param group, processed as one batched tensor.

    Work happens on the *physical* (contiguous) arrangement of the stacked tensor: conv weights
    kept channels_last by :class:`FlatParameters` are ``[n, O, kh, kw, I]`` in memory, and
    ``axis_map[logical_axis]`` gives the physical position of each logical axis so the axes are
    still preconditioned in the reference's logical order (0: O, 1: I, 2: kh, 3: kw)."""

    def __init__(self, params: List[torch.Tensor]):
        self.params = params
        self.shape = tuple(params[0].shape)
        self.factors: Dict[int, FactorBatch] = {}
        self.momentum: Optional[torch.Tensor] = None
        self._flat = None                       # (grad_view, param_view): logical [n, *shape] views
        self.flat_owner = None                  # FlatParameters backing this group (fused update path)
        self.flat_idxs: List[int] = []
        nd = len(self.shape)
        self.axis_map = list(range(nd))         # logical axis -> physical axis (without the batch dim)
        self._to_phys = None                    # permutation logical[n,*shape] -> physical contiguous

    def bind_flat(self):
        """Use zero-copy ``[n, *shape]`` views when the params sit contiguously in a flat buffer."""
        g0 = getattr(self.params[0], "_fdt_flat", None)
        if g0 is None:
            return
        flat = g0[0]
        idxs = []
        for p in self.params:
            f = getattr(p, "_fdt_flat", None)
            if f is None or f[0] is not flat:
                return
            idxs.append(f[1])
        views = flat.stacked_views(idxs, self.shape)
        if views is None:
            return
        self._flat = views
        self.flat_owner, self.flat_idxs = flat, idxs
        if len(self.shape) == 4 and not views[0].is_contiguous():
            # physical [n, O, kh, kw, I]
            self._to_phys = (0, 1, 3, 4, 2)
            self.axis_map = [0, 3, 1, 2]

    def _phys(self, t: torch.Tensor) -> torch.Tensor:
        return t if self._to_phys is None else t.permute(*self._to_phys)

    def grads(self) -> torch.Tensor:
        if self._flat is not None:
            return self._phys(self._flat[0])
        return torch.stack([p.grad for p in self.params])

    def values(self) -> torch.Tensor:
        if self._flat is not None:
            return self._phys(self._flat[1])
        return torch.stack([p.data for p in self.params])

    def write_back(self, new_values: torch.Tensor):
        if self._flat is not None:
            return  # updated in place through the view
        for p, v in zip(self.params, new_values.unbind(0)):
            p.data.copy_(v)


class NGD(Optimizer):
    """Drop-in for the reference's ``NGD(params, lr, momentum, dampening, weight_decay, nesterov,
    ngd, alpha, rank, update_period, eta)`` (ngd_optimizer.py:423-450)."""

    def __init__(self, params, lr=1e-4, momentum=0, dampening=0, weight_decay=0, nesterov=False,
                 ngd=True, alpha=4, rank=-1, update_period=4, eta=0.1):
        if lr < 0.0:
            raise ValueError(f"Invalid learning rate: {lr}")
        if momentum < 0.0:
            raise ValueError(f"Invalid momentum value: {momentum}")
        if weight_decay < 0.0:
            raise ValueError(f"Invalid weight_decay value: {weight_decay}")
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                        nesterov=nesterov, ngd=ngd, alpha=alpha, rank=rank,
                        update_period=update_period, eta=eta)
        super().__init__(params, defaults)
        self._groups: Dict[int, List[_ShapeGroup]] = {}
        self._lr_tensors: Dict[int, torch.Tensor] = {}
        self._hyper_host: Dict[int, Tuple[float, float, float]] = {}
        self._wd_done = False
        self.capturable = False

    # ------------------------------------------------------------------ grouping
    def _shape_groups(self, gi: int, group) -> List[_ShapeGroup]:
        if gi in self._groups:
            return self._groups[gi]
        buckets: "OrderedDict[Tuple, List[torch.Tensor]]" = OrderedDict()
        for p in group["params"]:
            if p.grad is None:
                continue
            f = getattr(p, "_fdt_flat", None)
            key = (tuple(p.shape), p.dtype, p.device, f[0].group_id_of(p) if f is not None else None)
            buckets.setdefault(key, []).append(p)
        sgs = [_ShapeGroup(ps) for ps in buckets.values()]
        for sg in sgs:
            sg.bind_flat()
        self._groups[gi] = sgs
        return sgs

    # ------------------------------------------------------------------ CUDA-graph support
    def _factor_batches(self):
        for sgs in self._groups.values():
            for sg in sgs:
                yield from sg.factors.values()

    def next_step_kind(self) -> str:
        """'warm' (first call: factors are created and warm-started), 'update' (Fisher refresh)
        or 'plain'.  The kind decides which kernels a step launches, i.e. which captured graph
        can replay it."""
        if not any(g["ngd"] for g in self.param_groups):
            return "plain" if self._groups else "warm"      # plain SGD: one eager step builds the groups
        fbs = list(self._factor_batches())
        if not fbs or any(fb.t == 0 for fb in fbs):
            return "warm"
        kinds = {fb.updating() for fb in fbs}
        if any(not g["ngd"] for g in self.param_groups) and not kinds:
            return "plain"
        return "update" if True in kinds else "plain"

    def advance_without_python(self):
        """Account for a step that ran as a CUDA-graph replay (no Python executed)."""
        for fb in self._factor_batches():
            fb.t += 1

    def set_capturable(self, flag: bool = True):
        """Keep lr / momentum / dampening in device scalars so a captured CUDA graph follows their schedules
        (MultiStepLR, OneCycleLR with ``cycle_momentum``) without re-capture."""
        self.capturable = flag

    @staticmethod
    def _hyper_values(group) -> Tuple[float, float, float]:
        return (float(group["lr"]), float(group["momentum"]), float(group["dampening"]))

    def _hyper(self, gi: int, group, device) -> torch.Tensor:
        """Device triple {lr, momentum, dampening} of a param group, created WITH the current values (a zero-initialised
        scalar would turn the first step after construction / ``load_state_dict`` into a no-op)."""
        t = self._lr_tensors.get(gi)
        if t is None or t.device != device:
            vals = self._hyper_values(group)
            t = torch.tensor(vals, dtype=torch.float32, device=device)
            self._lr_tensors[gi] = t
            self._hyper_host[gi] = vals
        return t

    def _lr(self, gi: int, group, device):
        if not self.capturable:
            return group["lr"]
        return self._hyper(gi, group, device)[0]

    def sync_lr(self):
        """Copy each group's python ``lr`` / ``momentum`` / ``dampening`` into its device triple when they changed
        (call outside graph replay; a no-op while a stream is capturing)."""
        for gi, group in enumerate(self.param_groups):
            t = self._lr_tensors.get(gi)
            if not torch.is_tensor(t):
                continue
            vals = self._hyper_values(group)
            if self._hyper_host.get(gi) != vals:
                if t.is_cuda and torch.cuda.is_current_stream_capturing():
                    continue
                t.copy_(torch.tensor(vals, dtype=torch.float32))
                self._hyper_host[gi] = vals

    def _begin_sums_arena(self):
        """The Σx² / Σx̂² accumulators of every factor batch live in ONE buffer cleared by one fill per step (they were
        ~60 separate zero-fills).  Built once every factor batch exists (after the first step); before that each call
        allocates its own."""
        arena = getattr(self, "_sums_arena", None)
        if arena is not None:
            arena.zero_()
            return
        fbs = [fb for fb in self._factor_batches() if fb.W.is_cuda and fb.W.dtype == torch.float32]
        if not fbs or any(fb.t == 0 for fb in fbs):
            return
        total = sum(2 * fb.n for fb in fbs)
        arena = torch.zeros(total, dtype=torch.float32, device=fbs[0].W.device)
        off = 0
        for fb in fbs:
            fb._sums = arena[off:off + 2 * fb.n]
            off += 2 * fb.n
        self._sums_arena = arena

    def fused_clip_wd(self, sq: torch.Tensor, max_norm: Optional[float], skipped: Optional[torch.Tensor]) -> bool:
        """Global-norm clip + non-finite guard + weight decay over the flat gradient range in ONE kernel,
        called by the Trainer between the gradient reduction and :meth:`step`.  Returns False (nothing done) when the
        parameters are not flat-backed or not on the native path; the caller then clips separately."""
        spans = []
        for gi, group in enumerate(self.param_groups):
            sgs = self._shape_groups(gi, group)
            span = self._flat_span(gi, sgs)
            if span is None:
                if any(sg.flat_owner is not None for sg in sgs):
                    return False
                continue
            spans.append((span, float(group["weight_decay"])))
        if not spans or not spans[0][0][0].grad_flat.is_cuda:
            return False
        from .. import ops
        if not ops.use_native(spans[0][0][0].grad_flat):
            return False
        for (flat, lo, hi), wd in spans:
            ops._cu().clip_wd_(flat.grad_flat[lo:hi], flat.param_flat[lo:hi], sq, max_norm, wd, skipped)
            skipped = None                      # count a skipped step once
        self._wd_done = True
        return True

    # ------------------------------------------------------------------ step
    def _precondition_group(self, sg: _ShapeGroup, group, d_p: torch.Tensor, pending=None) -> torch.Tensor:
        """Sequentially precondition every (logical) axis of the stacked physical tensor."""
        n = d_p.shape[0]
        nd = len(sg.shape)
        phys_shape = tuple(d_p.shape[1:])
        chain = self._norm_chain(sg, d_p)
        for axis in range(nd):
            D = sg.shape[axis]
            if D == 1:
                continue
            pa = sg.axis_map[axis]
            N = d_p[0].numel() // D
            fb = sg.factors.get(axis)
            if fb is None:
                fb = sg.factors[axis] = FactorBatch(n, N, D, group["alpha"], group["rank"], group["update_period"],
                                                    group["eta"], d_p.dtype, d_p.device)
            if _small_axis_native(fb, d_p):
                # short axis (conv kernel axes, class axes): one bandwidth-bound pass over the tensor where it lies
                d_p = precondition_small_axis(fb, d_p, pa, pending, chain)
            elif pa == nd - 1:
                d_p = precondition(fb, d_p.reshape(n, N, D), False, pending, chain).reshape(n, *phys_shape)
            elif pa == 0:
                d_p = precondition(fb, d_p.reshape(n, D, N), True, pending, chain).reshape(n, *phys_shape)
            else:
                Xm = d_p.movedim(pa + 1, -1)
                mid = Xm.shape
                d_p = precondition(fb, Xm.reshape(n, N, D), False, pending, chain).reshape(mid).movedim(-1, pa + 1)
        if chain is not None:
            d_p = chain.finish(d_p)
        return d_p

    def _norm_chain(self, sg, d_p):
        """A :class:`fdt.ops.cuda_ops.NormChain` when every axis of this tensor stack runs on a native kernel that can leave
        its norm restoration to the end of the chain and no factor refreshes its Fisher estimate on this step."""
        if not (d_p.is_cuda and d_p.dtype == torch.float32) or os.environ.get("FDT_NGD_NORM_CHAIN", "1") != "1":
            return None
        from .. import ops
        if not ops.use_native(d_p):
            return None
        cu = ops._cu()
        nd = len(sg.shape)
        links = 0
        for axis in range(nd):
            D = sg.shape[axis]
            if D == 1:
                continue
            fb = sg.factors.get(axis)
            if fb is None or fb.t == 0 or fb.updating():
                return None
            pa = sg.axis_map[axis]
            if cu.ngd_small_axis_ok(fb.D, fb.R):
                links += 1
            elif pa in (0, nd - 1) and cu.ngd_tf32_ok(fb.n, fb.N, fb.D, fb.R, pa == 0 and pa != nd - 1):
                links += 1
            else:
                return None
        return cu.NormChain(d_p) if links >= 2 else None

    def _side_streams(self, sgs: List[_ShapeGroup]):
        import os
        k = int(os.environ.get("FDT_NGD_STREAMS", "16"))
        if k <= 1 or not sgs or not sgs[0].params[0].is_cuda:
            return None
        key = ("streams", sgs[0].params[0].device)
        st = self._lr_tensors.get(key)
        if st is None:
            st = self._lr_tensors[key] = [torch.cuda.Stream(sgs[0].params[0].device) for _ in range(k)]
        return st

    def _flat_span(self, gi: int, sgs: List[_ShapeGroup]):
        """(flat, lo, hi) when this param group's flat-backed shape groups share one buffer."""
        key = ("span", gi)
        if key in self._lr_tensors:
            return self._lr_tensors[key]
        owners = {id(sg.flat_owner): sg.flat_owner for sg in sgs if sg.flat_owner is not None}
        span = None
        if len(owners) == 1:
            flat = next(iter(owners.values()))
            idxs = [i for sg in sgs if sg.flat_owner is not None for i in sg.flat_idxs]
            lo, hi = flat.span_of(idxs)
            # only safe when no foreign parameter lives inside [lo, hi)
            mine = set(idxs)
            inside = [i for i in range(len(flat.params)) if lo <= flat.span_of([i])[0] < hi]
            if all(i in mine for i in inside):
                span = (flat, lo, hi)
        self._lr_tensors[key] = span
        return span

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        pending: List[FactorBatch] = []
        self.sync_lr()
        self._begin_sums_arena()
        wd_done, self._wd_done = self._wd_done, False
        for gi, group in enumerate(self.param_groups):
            wd, mom, damp = group["weight_decay"], group["momentum"], group["dampening"]
            sgs = self._shape_groups(gi, group)
            span = self._flat_span(gi, sgs)
            lr = None
            if span is not None:
                flat, lo, hi = span
                gflat, pflat = flat.grad_flat[lo:hi], flat.param_flat[lo:hi]
                if wd != 0 and not wd_done:
                    gflat.add_(pflat, alpha=wd)
            # Shape groups are independent of each other (each is a chain of small batched kernels that
            # cannot fill 148 SMs), so on CUDA they are spread over a few side streams: forked from and
            # joined back into the current stream, which also works inside CUDA-graph capture.
            streams = self._side_streams(sgs)
            used = []
            spending = {}                                  # per side stream: its factor batches awaiting a Fisher refresh
            if streams:
                main = torch.cuda.current_stream()
                fork = torch.cuda.Event()
                fork.record(main)
            for si, sg in enumerate(sgs):
                fused = span is not None and sg.flat_owner is not None
                if streams and fused and group["ngd"]:
                    st = streams[si % len(streams)]
                    if st not in used:
                        st.wait_event(fork)                # fork (once per stream and step)
                        used.append(st)
                    with torch.cuda.stream(st):
                        G = sg.grads()
                        d_p = self._precondition_group(sg, group, G, spending.setdefault(st, []))
                        if d_p is not G:
                            G.copy_(d_p)
                    continue
                G = sg.grads()
                P = sg.values()
                if wd != 0 and not fused:
                    G.add_(P, alpha=wd)
                d_p = self._precondition_group(sg, group, G, pending) if group["ngd"] else G
                if fused:
                    if d_p is not G:
                        G.copy_(d_p)                      # result lands in the flat gradient buffer
                    continue
                if mom != 0:
                    if sg.momentum is None:
                        sg.momentum = d_p.clone(memory_format=torch.contiguous_format)
                    else:
                        sg.momentum.mul_(mom).add_(d_p, alpha=1.0 - damp)
                    d_p = d_p.add(sg.momentum, alpha=mom) if group["nesterov"] else sg.momentum
                lr = self._lr(gi, group, P.device)
                if torch.is_tensor(lr):
                    P.sub_(d_p * lr)
                else:
                    P.add_(d_p, alpha=-lr)
                sg.write_back(P)
            late = []
            if used:
                for st in used:                            # join only the streams that took work
                    ev = torch.cuda.Event()
                    ev.record(st)
                    torch.cuda.current_stream().wait_event(ev)
                    if spending.get(st):
                        # the refresh (eigenproblems + W update) of this stream's factors starts right here, overlapping
                        # the other streams' projections and the parameter update instead of trailing the whole step
                        with torch.cuda.stream(st):
                            flush_refreshes(spending[st])
                        late.append(st)
            if span is not None:
                self._fused_update(gi, group, gflat, pflat)
            for st in late:
                ev = torch.cuda.Event()
                ev.record(st)
                torch.cuda.current_stream().wait_event(ev)
        flush_refreshes(pending)
        return loss

    def _fused_update(self, gi: int, group, gflat: torch.Tensor, pflat: torch.Tensor):
        """Momentum + parameter update over the whole flat range: one kernel on CUDA."""
        mom, damp, nesterov = group["momentum"], group["dampening"], group["nesterov"]
        key = ("mom", gi)
        buf = self._lr_tensors.get(key)
        first = buf is None
        if first and mom != 0:
            buf = self._lr_tensors[key] = torch.zeros_like(gflat)
        if pflat.is_cuda:
            from .. import ops
            if ops.use_native(pflat):
                # lr, momentum and dampening are read from device memory: graph replays follow their schedules
                shadow = None
                span = self._lr_tensors.get(("span", gi))
                if span is not None and not span[0].sharded and span[0].shadow_flat is not None and len(self.param_groups) == 1:
                    # unsharded buffers: the bf16 operand shadow of the updated parameters comes out of this same pass
                    flat, lo, hi = span
                    shadow = flat.shadow_flat[lo:hi]
                    flat.shadow_fresh = True
                ops._cu().sgd_momentum_dev_(pflat, gflat, buf, self._hyper(gi, group, pflat.device), nesterov, first, shadow)
                return
        lr = self._lr(gi, group, pflat.device)
        d_p = gflat
        if mom != 0:
            if first:
                buf.copy_(d_p)
            else:
                buf.mul_(mom).add_(d_p, alpha=1.0 - damp)
            d_p = d_p.add(buf, alpha=mom) if nesterov else buf
        if torch.is_tensor(lr):
            pflat.sub_(d_p * lr)
        else:
            pflat.add_(d_p, alpha=-lr)

    # ------------------------------------------------------------------ (de)serialisation
    def state_dict(self):
        """Tensor-only state: per param group, per shape group: momentum + per-axis W/d/rho/t."""
        groups = []
        for gi, group in enumerate(self.param_groups):
            entry = {k: v for k, v in group.items() if k != "params"}
            sgs = []
            for sg in self._groups.get(gi, []):
                sgs.append({
                    "shape": sg.shape, "n": len(sg.params),
                    "momentum": None if sg.momentum is None else sg.momentum.detach().clone(),
                    "factors": {axis: {"t": fb.t, **{k: v.detach().clone() for k, v in fb.tensors().items()}}
                                for axis, fb in sg.factors.items()},
                })
            entry["shape_groups"] = sgs
            fm = self._lr_tensors.get(("mom", gi))
            entry["flat_momentum"] = None if fm is None else fm.detach().clone()
            groups.append(entry)
        return {"format": "fdt-ngd-v1", "groups": groups}

    def load_state_dict(self, state):
        """Validate the whole state against this optimizer's parameters first, then install it — a mismatching state
        (another optimizer, another rank's FSDP shard) raises ``ValueError`` and leaves the optimizer untouched."""
        if state.get("format") != "fdt-ngd-v1":
            raise ValueError("not an fdt NGD state dict")
        if len(state["groups"]) != len(self.param_groups):
            raise ValueError("parameter groups changed")
        plan = []
        for gi, (group, entry) in enumerate(zip(self.param_groups, state["groups"])):
            sgs = None
            if entry["shape_groups"]:
                # shape groups need gradients to exist to be discovered; build them from params
                for p in group["params"]:
                    if p.grad is None:
                        p.grad = torch.zeros_like(p)
                sgs = self._shape_groups(gi, group)
                if len(sgs) != len(entry["shape_groups"]):
                    raise ValueError("parameter set changed")
                for sg, s in zip(sgs, entry["shape_groups"]):
                    if tuple(s["shape"]) != sg.shape or s["n"] != len(sg.params):
                        raise ValueError(f"shape group mismatch: saved {tuple(s['shape'])} x{s['n']}, have {sg.shape} x{len(sg.params)}")
                    for axis, f in s["factors"].items():
                        D = sg.shape[int(axis)]
                        if f["W"].shape[0] != len(sg.params) or f["W"].shape[2] != D:
                            raise ValueError("factor shape mismatch")
            fm = entry.get("flat_momentum")
            if fm is not None:
                span = self._flat_span(gi, sgs) if sgs is not None else None
                if span is not None and fm.numel() != span[2] - span[1]:
                    raise ValueError(f"flat momentum has {fm.numel()} elements, this rank's span has {span[2] - span[1]}")
            plan.append((gi, group, entry, sgs))
        self._sums_arena = None
        for gi, group, entry, sgs in plan:
            for k, v in entry.items():
                if k not in ("shape_groups", "flat_momentum"):
                    group[k] = v
            if entry.get("flat_momentum") is not None:
                dev = group["params"][0].device
                self._lr_tensors[("mom", gi)] = entry["flat_momentum"].to(dev)
            if sgs is None:
                continue                       # saved before the first step: nothing to restore
            for sg, s in zip(sgs, entry["shape_groups"]):
                dev, dt = sg.params[0].device, sg.params[0].dtype
                sg.momentum = None if s["momentum"] is None else s["momentum"].to(dev, dt)
                sg.factors = {}
                for axis, f in s["factors"].items():
                    D = sg.shape[int(axis)]
                    N = sg.params[0].numel() // D
                    fb = FactorBatch(len(sg.params), N, D, group["alpha"], group["rank"],
                                     group["update_period"], group["eta"], dt, dev)
                    fb.t = int(f["t"])
                    fb.W.copy_(f["W"].to(dev, dt))          # into the padded storage
                    fb.d, fb.rho = f["d"].to(dev, dt), f["rho"].to(dev, dt)
                    sg.factors[int(axis)] = fb
