import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fdt.ops import cuda_ops as cu
from fdt.ops._lib import call, i64
torch.manual_seed(0)
def rel(a, b): return ((a - b).norm() / b.norm()).item()
for (n, N, D, R) in [(2, 256, 64, 32), (3, 200, 64, 32), (2, 100, 160, 80), (1, 4608, 512, 80), (2, 512, 512, 80)]:
    for left in (0, 1):
        if left and N % 4: continue
        X = torch.randn(n, N, D, device="cuda")
        W = torch.randn(n, R, D, device="cuda") * 0.1
        Xin = X.transpose(1, 2).contiguous() if left else X
        sums = torch.zeros(2 * n, device="cuda")
        Wt = W.transpose(1, 2).contiguous()
        H = torch.empty(n, N, R, device="cuda"); Xh = torch.empty_like(Xin)
        call("fdt_ngd_project", X.contiguous(), W, Wt, H, Xin, Xh, sums, n, N, D, R, left)
        Href = torch.bmm(X, W.transpose(1, 2))
        Xhref = X - torch.bmm(Href, W)
        Xh_cmp = Xh.transpose(1, 2) if left else Xh
        J = torch.zeros(n, R, D, device="cuda")
        cu.ngd_stats_j(Href.contiguous(), Xin, J, bool(left))
        Jref = torch.bmm(Href.transpose(1, 2), X)
        print(f"n{n} N{N} D{D} R{R} left{left}: H {rel(H, Href):.2e}  Xhat {rel(Xh_cmp, Xhref):.2e}  p0 {rel(sums[:n], (X*X).sum((1,2))):.1e} p1 {rel(sums[n:], (Xhref*Xhref).sum((1,2))):.1e}  J {rel(J, Jref):.2e}", flush=True)
