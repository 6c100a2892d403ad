"""Python face of ``csrc/gemm_sm100.cu`` / ``csrc/wgrad_sm100.cu``: the tcgen05 / TMEM / TMA GEMM,
implicit-GEMM convolution and weight-gradient kernels.

Every function returns ``None`` when the shape is outside what the native kernels cover, so the
caller (``cuda_ops``) can fall back to the library path for that call only.  Coverage:
K and N multiples of 64; convolutions with stride 1, "same" padding, Cin/Cout multiples of 64 and a
feature-map width that is a power of two ≤ 128 — i.e. every FusedConvBN layer of the CIFAR ResNets
except the 3-channel stem, and every Linear of the transformer.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import _lib
from ._lib import call, i64

BF16 = torch.bfloat16
ACT = {"none": 0, "relu": 1, "celu": 2, "gelu": 3, "tanh": 4}
_DISABLED = os.environ.get("FDT_GEMM", "auto") == "library"
_GEMM2 = os.environ.get("FDT_GEMM2", "1") == "1"      # route large GEMMs to the 2-CTA kernel (csrc/gemm2_sm100.cu)


def available() -> bool:
    if _DISABLED:
        return False
    try:
        return hasattr(_lib.lib(), "fdt_gemm_nt")
    except Exception:
        return False


def gemm_nt(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, act: str = "none",
            out_dtype=BF16, sums: Optional[torch.Tensor] = None, addend: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """act(a[M,K] @ b[N,K]ᵀ + bias) [+ addend] with fp32 accumulation; a, b bf16 row-major; ``addend`` bf16 [M, N]."""
    M, K = a.shape
    N = b.shape[0]
    if K % 64 or N % 64 or M < 1 or not (a.is_contiguous() and b.is_contiguous()):
        return None
    out = torch.empty(M, N, dtype=out_dtype, device=a.device)
    if _GEMM2 and sums is None and out_dtype == BF16 and N % 256 == 0 and M >= 1024:
        # large plain GEMMs (the transformer's Linear layers): 2-CTA clusters, tcgen05.mma.cta_group::2 on 256 x 256 tiles
        call("fdt_gemm2_nt", a, b, out, bias, addend, M, N, K, ACT[act])
        return out
    o16 = out if out_dtype == BF16 else None
    o32 = out if out_dtype == torch.float32 else None
    call("fdt_gemm_nt_ex", a, b, o16, o32, bias, sums, addend, M, N, K, ACT[act])
    return out


def gemm2_nt(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, act: str = "none",
             addend: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """The 2-CTA kernel (csrc/gemm2_sm100.cu): clusters of two CTAs, ``tcgen05.mma.cta_group::2`` on 256 x 256 tiles.
    bf16 out; N % 256 == 0, K % 64 == 0."""
    M, K = a.shape
    N = b.shape[0]
    if K % 64 or N % 256 or M < 1 or not (a.is_contiguous() and b.is_contiguous()):
        return None
    out = torch.empty(M, N, dtype=BF16, device=a.device)
    call("fdt_gemm2_nt", a, b, out, bias, addend, M, N, K, ACT[act])
    return out


def _conv_ok(x, w16, stride, padding) -> bool:
    """Shapes the implicit-GEMM kernel covers: stride 1 or 2, "same"-style padding (pad = k // 2), channel counts that are
    multiples of 64 and an OUTPUT width that is a power of two <= 128 (a 128-pixel tile = whole rows / whole images)."""
    N, Cin, H, W = x.shape
    Cout, _, kh, kw = w16.shape
    if stride not in (1, 2) or kh != kw or kh > 3 or padding != kh // 2 or Cin % 64 or Cout % 64:
        return False
    if stride == 2 and (H % 2 or W % 2):
        return False
    H, W = (H + 2 * padding - kh) // stride + 1, (W + 2 * padding - kw) // stride + 1
    if W < 1 or W > 128 or W & (W - 1) or 128 % W:
        return False
    bh = 128 // W
    if bh > H:
        return bh % H == 0
    return H % bh == 0


def conv_fwd(x: torch.Tensor, w16: torch.Tensor, stride: int, padding: int,
             sums: Optional[torch.Tensor] = None, addend: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """x: bf16 logical [N,Cin,H,W] channels_last; w16: bf16 logical [Cout,Cin,kh,kw] channels_last.  Stride 2 runs on
    the same implicit-GEMM kernel (the TMA tensor map walks every second pixel).  ``addend`` (bf16, shaped like the
    result) is added in the epilogue."""
    if not _conv_ok(x, w16, stride, padding):
        return None
    N, Cin, H, W = x.shape
    Cout, _, kh, _ = w16.shape
    Ho, Wo = (H + 2 * padding - kh) // stride + 1, (W + 2 * padding - kh) // stride + 1
    y = torch.empty((N, Cout, Ho, Wo), dtype=BF16, device=x.device, memory_format=torch.channels_last)
    if kh == 1 and stride == 1:
        call("fdt_gemm_nt_ex", x, w16, y, None, None, sums, addend, N * H * W, Cout, Cin, 0)
    else:
        call("fdt_conv_ex", x, w16, y, sums, addend, N, H, W, Cin, Cout, kh, padding, stride)
    return y


def dgrad_weight(w16: torch.Tensor) -> torch.Tensor:
    """[Cout,Cin,kh,kw] → the data-gradient filter [Cin,Cout,kh,kw] with taps rotated by 180°,
    channels_last (physically [Cin][kh][kw][Cout])."""
    return w16.flip(2, 3).transpose(0, 1).contiguous(memory_format=torch.channels_last)


def conv_dgrad(dy: torch.Tensor, w16: torch.Tensor, x_shape, stride: int, padding: int,
               wt: Optional[torch.Tensor] = None, addend: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """dX of a "same"-padded convolution on the implicit-GEMM kernel with the rotated, transposed filter (``wt``: that
    filter if the caller already has it — FlatParameters.refresh_dgrad_shadow).  Stride 1: one launch.  Stride 2: four
    parity-class launches that store straight to their pixels (``fdt_conv_dgrad_s2``) — no col2im.
    ``addend``: the gradient arriving at the same tensor through the other branch of a residual fork (bf16, shape of dX);
    it is added in the GEMM epilogue (stride 2: accumulated IN PLACE, the returned tensor is ``addend`` itself)."""
    N, Cin, H, W = x_shape
    Cout, _, kh, kw = w16.shape
    if wt is None:
        wt = dgrad_weight(w16)
    if stride == 1:
        if not _conv_ok(dy, wt, 1, padding):
            return None
        return conv_fwd(dy, wt, 1, padding, None, addend)
    if stride != 2 or kh != kw or kh > 3 or padding != kh // 2 or Cin % 64 or Cout % 64 or H % 2 or W % 2:
        return None
    Ho, Wo = H // 2, W // 2
    if tuple(dy.shape[2:]) != (Ho, Wo) or Wo > 128 or Wo & (Wo - 1):
        return None
    bh = 128 // Wo
    if (bh > Ho and bh % Ho) or (bh <= Ho and Ho % bh):
        return None
    if addend is not None:
        if not (addend.dtype == BF16 and addend.is_contiguous(memory_format=torch.channels_last)):
            addend = addend.to(dtype=BF16, memory_format=torch.channels_last)
        dx = addend
    else:
        dx = torch.empty((N, Cin, H, W), dtype=BF16, device=dy.device, memory_format=torch.channels_last)
    call("fdt_conv_dgrad_s2", dy, wt, dx, int(addend is not None), N, H, W, Cin, Cout, kh, padding, launches=4 if kh > 1 else 1)
    return dx


def conv_wgrad(dy, x, w_shape, stride, padding, accumulate_into=None, alpha=1.0, multicast=False):
    from . import wgrad_sm100
    return wgrad_sm100.conv_wgrad(dy, x, w_shape, stride, padding, accumulate_into, alpha, multicast)


# ----------------------------------------------------------------------------- explicit im2col path
def _im2col_geom(x_shape, w_shape, stride, padding):
    N, C, H, W = x_shape
    O, _, kh, kw = w_shape
    if kh != kw or O % 64:
        return None
    Ho = (H + 2 * padding - kh) // stride + 1
    Wo = (W + 2 * padding - kw) // stride + 1
    K = kh * kw * C
    Kp = K if C % 8 == 0 else (K + 63) // 64 * 64
    if Kp % 64:
        return None
    return N, C, H, W, O, kh, Ho, Wo, K, Kp


def _weight_2d(w16: torch.Tensor, K: int, Kp: int) -> torch.Tensor:
    """[O, I, kh, kw] channels_last → the [O, Kp] K-major GEMM operand (zero-padded columns for the stem)."""
    O = w16.shape[0]
    w2 = w16.permute(0, 2, 3, 1).reshape(O, K)
    if Kp != K:
        w2 = torch.nn.functional.pad(w2, (0, Kp - K))
    return w2.contiguous()


def im2col(x: torch.Tensor, geom, stride, padding) -> torch.Tensor:
    N, C, H, W, O, k, Ho, Wo, K, Kp = geom
    A = torch.empty(N * Ho * Wo, Kp, dtype=BF16, device=x.device)
    call("fdt_im2col", x, A, N, H, W, C, k, stride, padding, Ho, Wo, Kp)
    return A


def conv_fwd_im2col(x, w16, stride, padding, sums=None, keep=None):
    """``keep``: a list that receives the im2col matrix, so the weight-gradient GEMM of the backward pass can reuse
    it instead of rebuilding it."""
    geom = _im2col_geom(x.shape, w16.shape, stride, padding)
    if geom is None:
        return None
    N, C, H, W, O, k, Ho, Wo, K, Kp = geom
    A = im2col(x, geom, stride, padding)
    y = torch.empty((N, O, Ho, Wo), dtype=BF16, device=x.device, memory_format=torch.channels_last)
    call("fdt_gemm_nt", A, _weight_2d(w16, K, Kp), y, None, None, sums, N * Ho * Wo, O, Kp, 0)
    if keep is not None:
        keep.append(A)
    return y


def conv_dgrad_im2col(dy, w16, x_shape, stride, padding, wt=None):
    geom = _im2col_geom(x_shape, w16.shape, stride, padding)
    if geom is None:
        return None
    N, C, H, W, O, k, Ho, Wo, K, Kp = geom
    if wt is None or tuple(wt.shape) != (Kp, O):
        wt = _weight_2d(w16, K, Kp).t().contiguous()             # [Kp, O]
    dA = torch.empty(N * Ho * Wo, Kp, dtype=BF16, device=dy.device)
    call("fdt_gemm_nt", dy, wt, dA, None, None, None, N * Ho * Wo, Kp, O, 0)
    dx = torch.empty((N, C, H, W), dtype=BF16, device=dy.device, memory_format=torch.channels_last)
    call("fdt_col2im", dA, dx, N, H, W, C, k, stride, padding, Ho, Wo, Kp)
    return dx


def conv_wgrad_im2col(dy, x, w_shape, stride, padding, accumulate_into=None, alpha=1.0, multicast=False, A=None):
    from . import wgrad_sm100
    geom = _im2col_geom(x.shape, w_shape, stride, padding)
    if geom is None:
        return None
    N, C, H, W, O, k, Ho, Wo, K, Kp = geom
    if (N * Ho * Wo) % 64:
        return None
    if A is None or tuple(A.shape) != (N * Ho * Wo, Kp):
        A = im2col(x, geom, stride, padding)
    dy2 = dy.permute(0, 2, 3, 1).reshape(N * Ho * Wo, O)
    if accumulate_into is not None and Kp == K:
        return wgrad_sm100.linear_wgrad(dy2, A, accumulate_into, alpha, multicast)
    dw = wgrad_sm100.linear_wgrad(dy2, A)
    if dw is None:
        return None
    return dw[:, :K].reshape(O, k, k, C).permute(0, 3, 1, 2)


def linear(x, weight, bias, act, drop_p=0.0):
    from . import linear_sm100
    return linear_sm100.linear(x, weight, bias, act, drop_p)


def attention(q, k, v, key_mask, dropout_p, mask_compat):
    from . import attention_sm100
    return attention_sm100.attention(q, k, v, key_mask, dropout_p, mask_compat)
