#!/usr/bin/env python
"""Accuracy (vs fp64) and time of the 3xTF32 tcgen05 NGD products against the batched cuBLAS fp32 path."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fdt.ops import cuda_ops as cu
torch.manual_seed(0)
def rel(a, b): return ((a.double() - b).norm() / b.norm()).item()
def timeit(f, reps=20):
    for _ in range(3): f()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps): f()
    e.record(); e.synchronize()
    return s.elapsed_time(e) * 1e3 / reps
for (n, N, D, R) in [(2, 256, 64, 32), (3, 200, 64, 32), (2, 100, 160, 80), (3, 4608, 512, 80), (6, 2304, 256, 80), (3, 512, 2048, 80),
                     (4, 1152, 128, 64), (6, 1024, 256, 80), (12, 512, 512, 80), (1, 10, 2048, 80)]:
    for left in (0, 1):
        if left and N % 4: continue
        X = torch.randn(n, N, D, device="cuda")
        W = torch.randn(n, R, D, device="cuda") * 0.1
        Xin = X.transpose(1, 2).contiguous() if left else X
        H, out, p0 = cu.ngd_project(Xin, W, bool(left))
        Xd, Wd = X.double(), W.double()
        Href = torch.bmm(Xd, Wd.transpose(1, 2))
        Xhref = Xd - torch.bmm(Href, Wd)
        want = Xhref * torch.sqrt((Xd * Xd).sum((1, 2)) / (Xhref * Xhref).sum((1, 2)))[:, None, None]
        got = out.transpose(1, 2) if left else out
        J = torch.empty(n, R, D, device="cuda")
        Hf = Href.float().contiguous()
        cu.ngd_stats_j(Hf, Xin, J, bool(left))
        Jref = torch.bmm(Hf.double().transpose(1, 2), Xd)
        # fp32 cuBLAS reference error and time
        def torch_path():
            if left:
                Ht = torch.bmm(W, Xin); Xh = torch.baddbmm(Xin, W.transpose(1, 2), Ht, alpha=-1.0)
            else:
                Ht = torch.bmm(Xin, W.transpose(1, 2)); Xh = torch.baddbmm(Xin, Ht, W, alpha=-1.0)
            return cu.ngd_norms_rescale(Xin, Xh)
        t_tc = timeit(lambda: cu.ngd_project(Xin, W, bool(left)))
        t_bl = timeit(torch_path)
        t_j = timeit(lambda: cu.ngd_stats_j(Hf, Xin, J, bool(left)))
        t_jb = timeit(lambda: torch.bmm(Hf.transpose(1, 2), X))
        Hb = torch.bmm(X, W.transpose(1, 2))
        print(f"n{n} N{N} D{D} R{R} left{left}: H {rel(H, Href):.1e} (cuBLAS {rel(Hb, Href):.1e})  Xhat {rel(got, want):.1e}  J {rel(J, Jref):.1e} | "
              f"project {t_tc:6.1f} us vs {t_bl:6.1f} | J {t_j:6.1f} us vs {t_jb:6.1f}", flush=True)
