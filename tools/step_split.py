#!/usr/bin/env python
"""Where does the ResNet-50 step go?  Graph-replayed step time with (a) the full NGD optimizer, (b) plain
SGD-momentum on the same flat buffers (ngd=False), (c) forward+backward only."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from fdt.cli import ResNetStep
from fdt.engine.trainer import Trainer
from fdt.models.resnet import resnet50
from fdt.optim import NGD
from fdt.parallel.ddp import DataParallelEngine

dev = torch.device("cuda", 0)
loader = bench.make_host_batches("resnet50", 128, 4, 128, 0)
batches = [tuple(t.to(dev) for t in b) for b in loader]


def run(ngd, noopt=False):
    torch.manual_seed(0)
    eng = DataParallelEngine(resnet50(10).to(dev))
    opt = NGD(eng.optimizer_params(), lr=0.01, momentum=0.9, weight_decay=1e-4, ngd=ngd)
    if noopt:
        opt.step = lambda closure=None: None
    tr = Trainer(eng, opt, ResNetStep(0.99, None), clip_norm=10.0, use_cuda_graph=True)
    for i in range(20):
        tr.step(batches[i % 4])
    torch.cuda.synchronize()
    ts = []
    for i in range(16):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); tr.step(batches[i % 4]); e.record(); e.synchronize()
        ts.append(s.elapsed_time(e))
    return sum(ts) / len(ts), min(ts), max(ts)


for name, kw in [("full NGD", dict(ngd=True)), ("SGD-momentum (ngd=False)", dict(ngd=False)), ("fwd+bwd+clip only", dict(ngd=False, noopt=True))]:
    a, lo, hi = run(**kw)
    print(f"{name:28s}: mean {a:6.2f} ms  (min {lo:.2f}, max {hi:.2f})", flush=True)
