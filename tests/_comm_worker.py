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


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from fdt.parallel.symm import SymmetricHeap
    heap = SymmetricHeap.create(dev)
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
    # ---- engines: symmetric-memory path must track the NCCL path
    from fdt.cli import ResNetStep
    from fdt.engine.trainer import Trainer
    from fdt.models.resnet import resnet18
    from fdt.optim import NGD
    from fdt.parallel.comm import Communicator
    from fdt.parallel.ddp import DataParallelEngine
    from fdt.parallel.fsdp import FullyShardedEngine

    def run(engine_cls, use_symm):
        torch.manual_seed(0)
        comm = Communicator()
        alloc = None
        if use_symm:
            comm.symm = heap
            alloc = heap.alloc_f32
        eng = engine_cls(resnet18(10).to(dev), comm, allocator=alloc)
        opt = NGD(eng.optimizer_params(), lr=0.02, momentum=0.9, weight_decay=1e-4)
        tr = Trainer(eng, opt, ResNetStep(0.0, None), clip_norm=10.0)
        g = torch.Generator().manual_seed(rank)
        losses = []
        for i in range(4):
            x = torch.randint(0, 255, (32, 32, 32, 3), dtype=torch.uint8, generator=g).to(dev)
            y = torch.randint(0, 10, (32,), generator=g).to(dev)
            torch.manual_seed(100 + i)          # same augmentation / mixup draws in both runs
            loss, _ = tr.step((x, y))
            losses.append(float(loss))
        if hasattr(eng, "pre_forward"):
            eng._gathered = getattr(eng, "_gathered", True)
            eng.pre_forward()
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
    t = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    if t.item() != 1.0:
        print(f"[rank {rank}] FAILED", flush=True)
        sys.exit(1)
    if rank == 0:
        print("COMM WORKER OK", flush=True)


if __name__ == "__main__":
    main()
