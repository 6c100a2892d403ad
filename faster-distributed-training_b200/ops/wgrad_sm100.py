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


def conv_wgrad(dy, x, w_shape, stride, padding) -> Optional[torch.Tensor]:
    if not available():
        return None
    Cout, Cin, kh, kw = w_shape
    N, _, H, W = x.shape
    if stride != 1 or kh != kw or padding != kh // 2 or Cin % 64 or Cout % 64:
        return None
    if W > 64 or W & (W - 1) or 64 % W:
        return None
    bh = 64 // W
    if (bh > H and bh % H) or (bh <= H and H % bh):
        return None
    # fp32 gradient in the physical [Cout, kh, kw, Cin] layout, returned as the logical [Cout, Cin, kh, kw] view
    dw = torch.zeros((Cout, kh, kw, Cin), dtype=torch.float32, device=x.device)
    call("fdt_conv_wgrad", dy, x, dw, N, H, W, Cin, Cout, kh, padding)
    return dw.permute(0, 3, 1, 2)
