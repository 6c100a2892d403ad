"""Padded-pitch NGD kernels vs torch on one factor (debug aid)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fdt.ops import cuda_ops as cu
from fdt.optim.ngd import FactorBatch, _alloc_stats
torch.manual_seed(0)
for (D, N, left) in [(1002, 128, True), (1000, 128, True), (2050, 96, True), (30522, 512, True)]:
    fb = FactorBatch(1, N, D, 4, -1, 4, 0.1, torch.float32, torch.device("cuda"))
    fb.W.copy_(torch.randn(1, fb.R, D, device="cuda") * 0.05)
    _alloc_stats(fb)
    X = torch.randn(1, D, N, device="cuda")          # left layout [n, D, N]
    H, out, p0 = cu.ngd_project(X, fb.Wp, True, None, D)
    Xm = X.transpose(1, 2)                            # [n, N, D]
    Hr = Xm @ fb.W.transpose(1, 2)
    Xh = Xm - Hr @ fb.W
    p0r, p1r = (Xm * Xm).sum(), (Xh * Xh).sum()
    outr = (Xh * torch.sqrt(p0r / p1r)).transpose(1, 2)
    rel = lambda a, b: ((a - b).norm() / b.norm()).item()
    cu.ngd_stats_j(H, X, fb.Jp, True, D)
    Jr = Hr.transpose(1, 2) @ Xm
    cu.ngd_stats_kl(fb.Jp, fb.Wp, fb.K, fb.L)
    print(f"D={D} N={N}: H {rel(H, Hr):.2e} out {rel(out, outr):.2e} p0 {rel(p0, p0r.reshape(1)):.2e} J {rel(fb.J, Jr):.2e} "
          f"K {rel(fb.K, Jr @ Jr.transpose(1, 2)):.2e} L {rel(fb.L, Jr @ fb.W.transpose(1, 2)):.2e} pad(J) {fb.Jp[..., D:].abs().max().item() if fb.Dp > D else 0:.1e}")
