import subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
code = '''
import sys, math, torch
sys.path.insert(0, %r)
from fdt.ops import attention_sm100 as at
B,H,L,masked,bwd = %d,%d,%d,%d,%d
torch.manual_seed(0)
mk = lambda: torch.randn(B, L, H*64, device="cuda").bfloat16().view(B, L, H, 64).transpose(1, 2).requires_grad_(True)
q,k,v = mk(),mk(),mk()
km = None
if masked:
    lens = torch.randint(L//3, L+1, (B,), device="cuda")
    km = (torch.arange(L, device="cuda")[None,:] < lens[:,None]).long()
o = at.attention(q,k,v,km,0.0,False)
torch.cuda.synchronize(); print("fwd ok", flush=True)
if bwd:
    o.backward(torch.randn_like(o)); torch.cuda.synchronize(); print("bwd ok", flush=True)
'''
for cfg in [(4,8,128,1,1),(3,8,77,1,1),(2,8,64,1,1),(1,1,96,0,1),(1,8,128,0,1)]:
    try:
        r = subprocess.run([sys.executable, "-c", code % ((ROOT,)+cfg)], capture_output=True, text=True, timeout=25)
        print(cfg, r.stdout.strip().replace("\n", " | "), '\n'.join(l for l in r.stderr.splitlines() if 'fdt attention' in l), flush=True)
    except subprocess.TimeoutExpired as e:
        print(cfg, "TIMEOUT after:", (e.stdout or b"").decode().strip().replace("\n", " | "), flush=True)
