"""Weight-gradient kernels (MN-major tcgen05 GEMM with split-K) — see ``csrc/wgrad_sm100.cu``."""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib
from ._lib import call


def available() -> bool:
    try:
        return hasattr(_lib.lib(), "fdt_conv_wgrad")
    except Exception:
        return False


def _conv_ok(x_shape, w_shape, stride, padding) -> bool:
    Cout, Cin, kh, kw = w_shape
    N, _, H, W = x_shape
    if stride not in (1, 2) or kh != kw or Cin % 64 or Cout % 8:
        return False
    if stride == 1 and padding != kh // 2:
        return False
    if stride == 2 and (H % 2 or W % 2 or padding != kh // 2):
        return False
    H, W = (H + 2 * padding - kh) // stride + 1, (W + 2 * padding - kw) // stride + 1     # output grid
    if W < 1 or W > 64 or W & (W - 1) or 64 % W:
        return False
    bh = 64 // W
    if bh > H:
        # several images per 64-pixel k-block (4x4 / 2x2 maps): any batch size — the images a box reaches beyond the batch
        # are zero-filled by the TMA unit and contribute nothing
        return bh % H == 0
    return H % bh == 0 and (N * H * W) % 64 == 0


def conv_wgrad(dy, x, w_shape, stride, padding, accumulate_into=None, alpha: float = 1.0, multicast: bool = False):
    """fp32 weight gradient of a stride-1 or stride-2 "same"-padded convolution.  ``accumulate_into`` (a physically
    [Cout, kh, kw, Cin]-contiguous fp32 tensor, e.g. the parameter's slice of the flat gradient buffer)
    receives the result by atomic accumulation and ``True`` is returned; otherwise a fresh tensor in the
    logical [Cout, Cin, kh, kw] view comes back."""
    if not available() or not _conv_ok(x.shape, w_shape, stride, padding):
        return None
    Cout, Cin, kh, kw = w_shape
    N, _, H, W = x.shape
    if accumulate_into is not None:
        # a tensor, or a raw (possibly multicast) device address
        dst = accumulate_into if torch.is_tensor(accumulate_into) else _lib.ctypes.c_void_p(int(accumulate_into))
        call("fdt_conv_wgrad_ex", dy, x, dst, N, H, W, Cin, Cout, kh, padding, stride, float(alpha), int(multicast))
        return True
    dw = torch.zeros((Cout, kh, kw, Cin), dtype=torch.float32, device=x.device)
    call("fdt_conv_wgrad_ex", dy, x, dw, N, H, W, Cin, Cout, kh, padding, stride, 1.0, 0)
    return dw.permute(0, 3, 1, 2)


def linear_wgrad(gz: torch.Tensor, x16: torch.Tensor, accumulate_into=None, alpha: float = 1.0, multicast: bool = False):
    """dW[N, K] = gzᵀ[N, M] · x[M, K] in fp32 (gz, x bf16 row-major)."""
    if not available():
        return None
    M, N = gz.shape
    K = x16.shape[1]
    if K % 64 or N % 8 or M < 1 or not (gz.is_contiguous() and x16.is_contiguous()):
        return None
    if accumulate_into is not None:
        dst = accumulate_into if torch.is_tensor(accumulate_into) else _lib.ctypes.c_void_p(int(accumulate_into))
        call("fdt_linear_wgrad", gz, x16, dst, M, N, K, float(alpha), int(multicast))
        return True
    dw = torch.zeros(N, K, dtype=torch.float32, device=gz.device)
    call("fdt_linear_wgrad", gz, x16, dw, M, N, K, 1.0, 0)
    return dw
