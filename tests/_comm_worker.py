"""Multi-rank worker (launched by torchrun from test_comm_multigpu.py): fused NVLink collectives vs
torch.distributed (NCCL), then DDP / FSDP engine equivalence."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def check(name, a, b, rtol=1e-5, atol=1e-6):
    ok = torch.allclose(a, b, rtol=rtol, atol=atol)
    if not ok:
        print(f"[rank {dist.get_rank()}] MISMATCH {name}: max err {(a - b).abs().max().item():.3e}", flush=True)
    return ok


def check_collectives(heap, dev, rank, world):
    ok = True
    n = 4 * 1024 * 1024 + 512 * world
    n = n // (4 * world) * (4 * world)
    buf = heap.alloc_f32(n)
    has_mc = bool(heap._ctx(buf).mc)
    if rank == 0:
        print(f"world={world} multicast={'yes' if has_mc else 'no'} n={n}", flush=True)
    for use_mc in ([True, False] if has_mc else [False]):
        heap.use_multicast = use_mc
        for it in range(3):
            g = torch.Generator(device=dev).manual_seed(1000 * it + rank)
            src = torch.randn(n, device=dev, generator=g)
            # all-reduce
            buf.copy_(src)
            ref = src.clone()
            dist.all_reduce(ref)
            ref /= world
            heap.all_reduce_mean_(buf)
            ok &= check(f"all_reduce mc={use_mc} it={it}", buf, ref)
            ok &= check("sqnorm(all_reduce)", heap.last_sqnorm, (ref.double() ** 2).sum().float(), rtol=1e-4, atol=0)
            # reduce-scatter
            buf.copy_(src)
            region = n // world
            lo, hi = rank * region, (rank + 1) * region
            out = heap.reduce_scatter_mean_(buf, lo, hi)
            ok &= check(f"reduce_scatter mc={use_mc} it={it}", out, ref[lo:hi])
            ok &= check("sqnorm(reduce_scatter)", heap.last_sqnorm, (ref.double() ** 2).sum().float(), rtol=1e-4, atol=0)
            # all-gather + bf16 cast
            full = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(77 + it))
            buf.zero_()
            buf[lo:hi] = full[lo:hi]
            shadow = torch.empty(n, dtype=torch.bfloat16, device=dev)
            heap.all_gather_(buf, lo, hi, shadow=shadow)
            ok &= check(f"all_gather it={it}", buf, full, 0, 0)
            ok &= bool(torch.equal(shadow, full.bfloat16()))
    heap.use_multicast = True
    return ok


def check_sizes(heap, dev, rank, world):
    """1 KB ... 1 GB, including a length that is not a multiple of the 16-byte vector (the allocator pads it)."""
    ok = True
    sizes = [1024, 4096 + 4, 1 << 20, 94 * 1024 * 1024, 1 << 30]
    if os.environ.get("COMM_WORKER_SMALL") == "1":
        sizes = sizes[:-1]
    for nbytes in sizes:
        n = nbytes // 4
        buf = heap.alloc_f32(n)                      # view of a padded symmetric allocation
        full = heap.bufs[buf.data_ptr()][0]
        assert buf.numel() == n and full.numel() % (4 * world) == 0
        g = torch.Generator(device=dev).manual_seed(nbytes % 9973 + rank)
        src = torch.randn(n, device=dev, generator=g)
        buf.copy_(src)
        ref = src.clone()
        dist.all_reduce(ref)
        ref /= world
        heap.all_reduce_mean_(buf)
        ok &= check(f"all_reduce {nbytes} B", buf, ref)
        ok &= bool((full[n:] == 0).all())            # the padding stays zero
        ok &= check(f"sqnorm {nbytes} B", heap.last_sqnorm, (ref.double() ** 2).sum().float(), rtol=1e-4, atol=0)
        # reduce-scatter / all-gather over the padded buffer's equal regions
        np_ = full.numel()
        region = np_ // world
        lo, hi = rank * region, (rank + 1) * region
        full[:n].copy_(src)
        out = heap.reduce_scatter_mean_(full, lo, hi)
        refp = torch.zeros(np_, device=dev)
        refp[:n] = ref
        ok &= check(f"reduce_scatter {nbytes} B", out, refp[lo:hi])
        truth = torch.randn(np_, device=dev, generator=torch.Generator(device=dev).manual_seed(5))
        full.zero_()
        full[lo:hi] = truth[lo:hi]
        shadow = torch.empty(np_, dtype=torch.bfloat16, device=dev)
        heap.all_gather_(full, lo, hi, shadow=shadow)
        ok &= check(f"all_gather {nbytes} B", full, truth, 0, 0)
        ok &= bool(torch.equal(shadow, truth.bfloat16()))
        heap.free(buf)
        del buf, full, src, ref, refp, truth, shadow
        torch.cuda.empty_cache()
    if rank == 0:
        print(f"sizes {sizes}: {'ok' if ok else 'MISMATCH'}", flush=True)
    return ok


def check_ranges(heap, dev, rank, world):
    """Bucketed collectives: per-rank ranges of different (and zero) length, a signal channel of their own, Σg²
    accumulated over two buckets."""
    ok = True
    lens = [4 * (1000 * (r + 1) + 37 * r) for r in range(world)]
    lens[-1] = 0                                     # one rank owns nothing in this bucket
    starts = [sum(lens[:r]) for r in range(world)]
    half = sum(lens)
    n = 2 * half                                     # bucket A = first half, bucket B = the same layout shifted by `half`
    buf = heap.alloc_f32(n)
    g = torch.Generator(device=dev).manual_seed(4242 + rank)
    src = torch.randn(n, device=dev, generator=g)
    ref = src.clone()
    dist.all_reduce(ref)
    ref /= world
    buf.copy_(src)
    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for b in range(2):
            lo = b * half + starts[rank]
            heap.reduce_scatter_range_(buf, lo, lo + lens[rank], blocks=3, channel=1)
        sq = heap.take_sq()
    torch.cuda.current_stream().wait_stream(side)
    for b in range(2):
        lo = b * half + starts[rank]
        ok &= check(f"reduce_scatter_range bucket {b}", buf[lo:lo + lens[rank]], ref[lo:lo + lens[rank]])
    ok &= check("sqnorm(ranges)", sq, (ref.double() ** 2).sum().float(), rtol=1e-4, atol=0)
    # all-gather by range table
    truth = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(99))
    shadow = torch.zeros(n, dtype=torch.bfloat16, device=dev)
    buf.zero_()
    for b in range(2):
        lo = b * half + starts[rank]
        buf[lo:lo + lens[rank]] = truth[lo:lo + lens[rank]]
    for b in range(2):
        table = torch.tensor([v for r in range(world) for v in (b * half + starts[r], b * half + starts[r] + lens[r])],
                             dtype=torch.int64, device=dev)
        with torch.cuda.stream(side):
            heap.all_gather_range_(buf, table, shadow, blocks=5, channel=1)
    torch.cuda.current_stream().wait_stream(side)
    ok &= check("all_gather_range", buf, truth, 0, 0)
    ok &= bool(torch.equal(shadow, truth.bfloat16()))
    # two-phase broadcast of one owner's range (odd length in vectors, not a multiple of world or of the block count)
    owner = world - 1
    lo, hi = 4 * 13, 4 * 13 + 4 * 100003
    nb = heap.alloc_f32(hi + 64)
    truth2 = torch.randn(hi + 64, device=dev, generator=torch.Generator(device=dev).manual_seed(123))
    nb.fill_(-7.0)
    if rank == owner:
        nb[lo:hi] = truth2[lo:hi]
    shadow2 = torch.zeros(hi + 64, dtype=torch.bfloat16, device=dev)
    torch.cuda.synchronize()
    dist.barrier()
    with torch.cuda.stream(side):
        heap.bcast_range_(nb, owner, lo, hi, shadow2, blocks=7, channel=1)
    torch.cuda.current_stream().wait_stream(side)
    ok &= check("bcast_range", nb[lo:hi], truth2[lo:hi], 0, 0)
    ok &= bool(torch.equal(shadow2[lo:hi], truth2[lo:hi].bfloat16()))
    ok &= bool((nb[:lo] == -7.0).all()) and bool((nb[hi:] == -7.0).all())
    if rank == 0:
        print(f"range collectives: {'ok' if ok else 'MISMATCH'}", flush=True)
    return ok


def finish(ok, dev, rank):
    t = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    if t.item() != 1.0:
        print(f"[rank {rank}] FAILED", flush=True)
        sys.exit(1)
    if rank == 0:
        print("COMM WORKER OK", flush=True)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from fdt.parallel.symm import SymmetricHeap
    heap = SymmetricHeap.create(dev)
    ok = True
    mode = os.environ.get("COMM_WORKER_MODE", "all")          # all | collectives | engines
    if mode in ("all", "collectives"):
        ok &= check_collectives(heap, dev, rank, world)
        ok &= check_sizes(heap, dev, rank, world)
        ok &= check_ranges(heap, dev, rank, world)
    if mode == "collectives":
        return finish(ok, dev, rank)
    # ---- engines: symmetric-memory path must track the NCCL path
    from fdt.cli import ResNetStep
    from fdt.engine.trainer import Trainer
    from fdt.models.resnet import resnet18
    from fdt.optim import NGD
    from fdt.parallel.comm import Communicator
    from fdt.parallel.ddp import DataParallelEngine
    from fdt.parallel.fsdp import FullyShardedEngine

    def run(engine_cls, use_symm, steps=4, **kw):
        torch.manual_seed(0)
        from fdt.ops import cuda_ops
        cuda_ops.reset_seed()               # same Philox stream (augmentation, mixup draws) in every run
        comm = Communicator()
        alloc = None
        if use_symm:
            comm.symm = heap
            alloc = heap.alloc_f32
        eng = engine_cls(resnet18(10).to(dev), comm, allocator=alloc, **kw)
        opt = NGD(eng.optimizer_params(), lr=0.02, momentum=0.9, weight_decay=1e-4)
        tr = Trainer(eng, opt, ResNetStep(0.0, None), clip_norm=10.0)
        g = torch.Generator().manual_seed(rank)
        losses = []
        for i in range(steps):
            x = torch.randint(0, 255, (32, 32, 32, 3), dtype=torch.uint8, generator=g).to(dev)
            y = torch.randint(0, 10, (32,), generator=g).to(dev)
            torch.manual_seed(100 + i)          # same augmentation / mixup draws in both runs
            loss, _ = tr.step((x, y))
            losses.append(float(loss))
        with torch.no_grad():
            eng.pre_forward()                   # collective: gathers what the owners updated in the last step
        torch.cuda.synchronize()
        return losses, torch.cat([v.detach().float().flatten() for k, v in sorted(eng.module.state_dict().items())
                                  if v.dtype.is_floating_point])

    for cls in (DataParallelEngine, FullyShardedEngine):
        l_a, p_a = run(cls, True)
        l_b, p_b = run(cls, False)
        rel = ((p_a - p_b).norm() / p_b.norm()).item()
        good = rel < 2e-2 and all(abs(a - b) < 0.1 for a, b in zip(l_a, l_b))
        if rank == 0:
            print(f"{cls.__name__}: losses symm {['%.4f' % v for v in l_a]} nccl {['%.4f' % v for v in l_b]} param rel diff {rel:.2e}", flush=True)
        ok &= good
    # The default DDP step (bucketed reduce-scatter, sharded NGD, staged all-gather) against the replicated step (one
    # all-reduce, the full NGD step on every rank), 8 steps.
    # (a) on IDENTICAL synthetic gradients — no forward/backward, so no atomics-order noise is amplified by the bf16
    #     network: the reduce / shard / NGD / gather pipeline itself must agree to 1e-3 (it agrees far tighter);
    # (b) end to end, relative to the run-to-run noise floor of the replicated path itself (split-K wgrad atomics make
    #     two identical runs drift apart by ~1e-2 after a few NGD steps).
    if os.environ.get("FDT_DDP_SHARD", "1") == "1":
        def synthetic(**kw):
            torch.manual_seed(0)
            comm = Communicator()
            comm.symm = heap
            eng = DataParallelEngine(resnet18(10).to(dev), comm, allocator=heap.alloc_f32, **kw)
            opt = NGD(eng.optimizer_params(), lr=0.02, momentum=0.9, weight_decay=1e-4)
            tr = Trainer(eng, opt, ResNetStep(0.0, None), clip_norm=10.0)
            plist = list(eng.module.parameters())
            for step in range(8):
                eng.zero_grad()
                with torch.no_grad():
                    eng.pre_forward()
                    for i, p in enumerate(plist):
                        g = torch.Generator(device=dev).manual_seed(100000 * step + 100 * i + rank)
                        p.grad.copy_(torch.randn(p.shape, device=dev, generator=g) * 0.05)
                eng.reduce_gradients()
                tr._clip()
                opt.step()
                eng.post_step()
            with torch.no_grad():
                eng.pre_forward()
            torch.cuda.synchronize()
            return torch.cat([v.detach().float().flatten() for k, v in sorted(eng.module.state_dict().items())
                              if v.dtype.is_floating_point])
        p_s, p_r = synthetic(), synthetic(shard_optimizer=False)
        rel = ((p_s - p_r).norm() / p_r.norm()).item()
        if rank == 0:
            print(f"DDP sharded vs replicated on identical gradients, 8 steps: param rel diff {rel:.2e}", flush=True)
        ok &= rel < 1e-3
        l_s, p_s = run(DataParallelEngine, True, steps=8)
        l_r, p_r = run(DataParallelEngine, True, steps=8, shard_optimizer=False)
        l_r2, p_r2 = run(DataParallelEngine, True, steps=8, shard_optimizer=False)
        rel = ((p_s - p_r).norm() / p_r.norm()).item()
        floor = ((p_r2 - p_r).norm() / p_r.norm()).item()
        if rank == 0:
            print(f"DDP sharded+overlapped vs replicated end to end, 8 steps: param rel diff {rel:.2e} "
                  f"(replicated vs itself: {floor:.2e}); losses {['%.4f' % v for v in l_s]} vs {['%.4f' % v for v in l_r]}",
                  flush=True)
        ok &= rel < max(1e-3, 3.0 * floor) and all(abs(a - b) < 0.15 for a, b in zip(l_s, l_r))
    finish(ok, dev, rank)


if __name__ == "__main__":
    main()
