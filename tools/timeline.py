#!/usr/bin/env python
"""Kernel timeline of one CUDA-graph-replayed training step (guidance only, taken under the CUPTI
profiler): per-stream busy time, the optimizer phase, and the kernels on the critical tail."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--bs", type=int, default=None)
    ap.add_argument("--kinds", default="plain,update")
    ap.add_argument("--dump", default=None)
    ap.add_argument("--mode", default="ddp", choices=["ddp", "fsdp"])
    args = ap.parse_args()
    import bench
    rank, local, world = bench.dist_env()          # under torchrun: one rank per GPU, rank 0 reports
    torch.cuda.set_device(local)
    bench.init_dist(world)
    dev = torch.device("cuda", local)
    bs = args.bs or (128 if args.model == "resnet50" else 64)
    tr, _ = bench.build_fdt(args.model, bs, dev, world, True, args.mode)
    loader = bench.make_host_batches(args.model, bs, 4, 128, rank)
    batches = [tuple(t.to(dev) for t in b) for b in loader]
    for i in range(24):
        tr.step(batches[i % 4])
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    want = set(args.kinds.split(","))
    seen = set()
    for it in range(12):
        kind = tr.optimizer.next_step_kind()
        if kind in seen or kind not in want:
            tr.step(batches[it % 4])
            torch.cuda.synchronize()
            continue
        seen.add(kind)
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            tr.step(batches[it % 4])
            torch.cuda.synchronize()
        evs = [e for e in prof.profiler.kineto_results.events() if e.device_type() == torch.autograd.DeviceType.CUDA]
        evs = sorted(evs, key=lambda e: e.start_ns())
        if rank != 0:
            continue
        if not evs:
            print("no device events")
            continue
        t0 = evs[0].start_ns()
        end = max(e.start_ns() + e.duration_ns() for e in evs)
        print(f"===== kind={kind}: span {(end - t0) / 1e6:.3f} ms, {len(evs)} device activities")
        streams = {}
        for e in evs:
            streams.setdefault(e.device_resource_id(), []).append(e)
        for sid, es in sorted(streams.items(), key=lambda kv: -sum(x.duration_ns() for x in kv[1]))[:6]:
            print(f"  stream {sid}: {len(es)} kernels, busy {sum(x.duration_ns() for x in es) / 1e6:.3f} ms, "
                  f"[{(es[0].start_ns() - t0) / 1e6:.3f}, {(es[-1].start_ns() + es[-1].duration_ns() - t0) / 1e6:.3f}] ms")
        # coarse phases by kernel name
        def first(pred):
            for e in evs:
                if pred(e.name()):
                    return (e.start_ns() - t0) / 1e6
            return None
        def last(pred):
            r = None
            for e in evs:
                if pred(e.name()):
                    r = (e.start_ns() + e.duration_ns() - t0) / 1e6
            return r
        print("  first wgrad at", first(lambda n: "wgrad" in n), " last wgrad end", last(lambda n: "wgrad" in n))
        print("  first ngd kernel at", first(lambda n: "ngd" in n), " last", last(lambda n: "ngd" in n))
        print("  sgd_momentum at", first(lambda n: "sgd_momentum" in n))
        if world > 1:
            # collectives of the sharded step: reduce-scatter buckets (should start inside backward, i.e. before the last
            # wgrad ends) and the staged all-gather / broadcast (should overlap the forward pass)
            for e in evs:
                n = e.name()
                if "reduce_kernel" in n or "all_gather" in n or "bcast_two_phase" in n or "take_scalar" in n:
                    print(f"  comm  {(e.start_ns() - t0) / 1e6:8.3f} ms  +{e.duration_ns() / 1e3:7.1f} us  stream {e.device_resource_id():<4d} {n[:70]}")
        # gaps on the union timeline (no kernel running anywhere)
        cur = t0
        idle = 0
        for e in evs:
            if e.start_ns() > cur:
                idle += e.start_ns() - cur
            cur = max(cur, e.start_ns() + e.duration_ns())
        print(f"  device idle (no kernel on any stream): {idle / 1e6:.3f} ms")
        if args.dump:
            with open(f"{args.dump}_{kind}.txt", "w") as f:
                for e in evs:
                    f.write(f"{(e.start_ns() - t0) / 1e3:10.1f} us  {e.duration_ns() / 1e3:8.1f} us  s{e.device_resource_id():<4d} {e.name()[:120]}\n")


if __name__ == "__main__":
    main()
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()
