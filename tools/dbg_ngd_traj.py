"""One-tensor NGD trajectory: native vs torch path vs the reference (debug aid)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from _refmods import load
from fdt import ops
from fdt.optim.ngd import NGD
from fdt.parallel.flat import FlatParameters
ref = load("ngd_optimizer")
shape = tuple(int(v) for v in sys.argv[1].split(","))
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
g = torch.Generator().manual_seed(0)
p0 = (torch.randn(*shape, generator=g) * 0.1).cuda()
def run(kind):
    p = torch.nn.Parameter(p0.clone())
    if kind == "ref":
        o = ref.NGD([p], lr=0.01, momentum=0.9, weight_decay=1e-4)
    else:
        FlatParameters([p], torch.device("cuda"))
        ops.force_torch(kind == "torch")
        o = NGD([p], lr=0.01, momentum=0.9, weight_decay=1e-4)
    gg = torch.Generator(device="cuda").manual_seed(1)
    traj = []
    for s in range(steps):
        grad = torch.randn(shape, device="cuda", generator=gg)
        if kind == "ref":
            p.grad = grad.clone()
        else:
            p.grad.copy_(grad)
        o.step()
        traj.append(p.detach().clone())
    ops.force_torch(False)
    return traj
r, n, t = run("ref"), run("native"), run("torch")
for s in range(steps):
    f = lambda a, b: ((a - b).norm() / b.norm()).item()
    print(f"step {s}: native-ref {f(n[s], r[s]):.2e}  torch-ref {f(t[s], r[s]):.2e}  native-torch {f(n[s], t[s]):.2e}  finite {torch.isfinite(n[s]).all().item()}")
