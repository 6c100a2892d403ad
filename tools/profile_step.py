#!/usr/bin/env python
"""Kernel-time breakdown of one training step (guidance only — numbers taken under a profiler are
never reported as benchmark values).  Prints the top CUDA kernels of a 'plain' and of a
'Fisher-refresh' NGD step of the flagship models in eager mode."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--bs", type=int, default=128)
    ap.add_argument("--top", type=int, default=25)
    args = ap.parse_args()
    import bench
    dev = torch.device("cuda", 0)
    tr, _ = bench.build_fdt(args.model, args.bs, dev, 1, False, "ddp")
    loader = bench.make_host_batches(args.model, args.bs, 4, 128, 0)
    batches = [tuple(t.to(dev) for t in b) for b in loader]
    for i in range(13):
        tr.step(batches[i % 4])
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    for name in ("first", "second", "third", "fourth"):
        kind = tr.optimizer.next_step_kind()
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            tr.step(batches[0])
            torch.cuda.synchronize()
        ev = prof.key_averages()
        tot = sum(e.device_time_total for e in ev if e.device_type == torch.autograd.DeviceType.CUDA)
        print(f"===== step kind={kind}: total device kernel time {tot/1e3:.2f} ms, {sum(e.count for e in ev if e.device_type == torch.autograd.DeviceType.CUDA)} kernels")
        if kind in ("update",) or name == "second":
            rows = sorted([e for e in ev if e.device_type == torch.autograd.DeviceType.CUDA], key=lambda e: -e.device_time_total)
            for e in rows[: args.top]:
                print(f"{e.device_time_total/1e3:9.3f} ms  x{e.count:<5d} {e.key[:110]}")


if __name__ == "__main__":
    main()
