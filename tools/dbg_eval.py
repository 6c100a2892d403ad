import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fdt.models.resnet import resnet50
import copy
torch.manual_seed(0)
for mode in ("train", "eval"):
    m = resnet50(10); g = copy.deepcopy(m).cuda()
    (m.train() if mode == "train" else m.eval()); (g.train() if mode == "train" else g.eval())
    x = torch.randn(48, 3, 32, 32)
    outs_c, outs_g = {}, {}
    def hook(store, name):
        def f(mod, inp, out): store[name] = out.detach().float().cpu()
        return f
    for (n1, b1), (n2, b2) in zip(m.named_modules(), g.named_modules()):
        if n1.count(".") == 1 and n1.split(".")[0].startswith("conv") and n1.split(".")[1].isdigit() and not n1.startswith("conv1"):
            b1.register_forward_hook(hook(outs_c, n1)); b2.register_forward_hook(hook(outs_g, n2))
    with torch.no_grad():
        rc = m(x); rg = g(x.cuda()).cpu()
    print(mode, "final rel err", ((rg - rc).norm() / rc.norm()).item())
    for k in outs_c:
        a, b = outs_g[k], outs_c[k]
        print(f"   {k:12s} rel {((a-b).norm()/b.norm()).item():.4f}  ref absmax {b.abs().max().item():.3g}")
