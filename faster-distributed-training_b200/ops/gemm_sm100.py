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


def available() -> bool:
    if _DISABLED:
        return False
    try:
        return hasattr(_lib.lib(), "fdt_gemm_nt")
    except Exception:
        return False


def gemm_nt(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, act: str = "none",
            out_dtype=BF16, sums: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """act(a[M,K] @ b[N,K]ᵀ + bias) with fp32 accumulation; a, b bf16 row-major."""
    M, K = a.shape
    N = b.shape[0]
    if K % 64 or N % 64 or M < 1 or not (a.is_contiguous() and b.is_contiguous()):
        return None
    out = torch.empty(M, N, dtype=out_dtype, device=a.device)
    o16 = out if out_dtype == BF16 else None
    o32 = out if out_dtype == torch.float32 else None
    call("fdt_gemm_nt", a, b, o16, o32, bias, sums, M, N, K, ACT[act])
    return out


def _conv_ok(x, w16, stride, padding) -> bool:
    N, Cin, H, W = x.shape
    Cout, _, kh, kw = w16.shape
    if stride != 1 or kh != kw or padding != kh // 2 or Cin % 64 or Cout % 64:
        return False
    if W > 128 or W & (W - 1) or 128 % W:
        return False
    bh = 128 // W
    if bh > H:
        return bh % H == 0
    return H % bh == 0


def conv_fwd(x: torch.Tensor, w16: torch.Tensor, stride: int, padding: int,
             sums: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """x: bf16 logical [N,Cin,H,W] channels_last; w16: bf16 logical [Cout,Cin,kh,kw] channels_last."""
    if not _conv_ok(x, w16, stride, padding):
        return None
    N, Cin, H, W = x.shape
    Cout, _, kh, _ = w16.shape
    y = torch.empty((N, Cout, H, W), dtype=BF16, device=x.device, memory_format=torch.channels_last)
    if kh == 1:
        call("fdt_gemm_nt", x, w16, y, None, None, sums, N * H * W, Cout, Cin, 0)
    else:
        call("fdt_conv_fwd", x, w16, y, sums, N, H, W, Cin, Cout, kh, padding)
    return y


def dgrad_weight(w16: torch.Tensor) -> torch.Tensor:
    """[Cout,Cin,kh,kw] → the data-gradient filter [Cin,Cout,kh,kw] with taps rotated by 180°,
    channels_last (physically [Cin][kh][kw][Cout])."""
    return w16.flip(2, 3).transpose(0, 1).contiguous(memory_format=torch.channels_last)


def conv_dgrad(dy: torch.Tensor, w16: torch.Tensor, x_shape, stride: int, padding: int) -> Optional[torch.Tensor]:
    """dX of a stride-1 "same" convolution = the same implicit GEMM with the rotated, transposed filter."""
    N, Cin, H, W = x_shape
    Cout, _, kh, kw = w16.shape
    wt = dgrad_weight(w16)
    if not _conv_ok(dy, wt, stride, padding):
        return None
    return conv_fwd(dy, wt, stride, padding)


def conv_wgrad(dy, x, w_shape, stride, padding, accumulate_into=None, alpha=1.0, multicast=False):
    from . import wgrad_sm100
    return wgrad_sm100.conv_wgrad(dy, x, w_shape, stride, padding, accumulate_into, alpha, multicast)


def linear(x, weight, bias, act):
    from . import linear_sm100
    return linear_sm100.linear(x, weight, bias, act)


def attention(q, k, v, key_mask, dropout_p, mask_compat):
    from . import attention_sm100
    return attention_sm100.attention(q, k, v, key_mask, dropout_p, mask_compat)
