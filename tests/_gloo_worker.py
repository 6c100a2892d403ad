"""2-process gloo worker: DDP and FSDP engines must reproduce the single-process trajectory on the
union batch (same model, SGD-free NGD optimizer, no mixup randomness)."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from fdt.engine.trainer import Trainer
    from fdt.optim import NGD
    from fdt.parallel import dist as pdist
    from fdt.parallel.comm import Communicator
    from fdt.parallel.ddp import DataParallelEngine
    from fdt.parallel.fsdp import FullyShardedEngine
    pdist.setup_norank(device="cpu")
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(0)

    def make_model():
        torch.manual_seed(0)
        # bias-free: 1-D parameters give rank-deficient NGD factors whose trajectory amplifies
        # summation-order noise (see tests/test_ngd.py), which would mask real engine errors
        return torch.nn.Sequential(torch.nn.Linear(12, 16, bias=False), torch.nn.Tanh(), torch.nn.Linear(16, 16, bias=False),
                                   torch.nn.Tanh(), torch.nn.Linear(16, 8, bias=False))   # every factor sees N >= R rows

    g = torch.Generator().manual_seed(5)
    X, Y = torch.randn(4, 16, 12, generator=g), torch.randint(0, 8, (4, 16), generator=g)

    def fwd_loss(eng, batch):
        x, y = batch
        out = eng(x)
        return torch.nn.functional.cross_entropy(out, y), out.detach().sum()

    def run(engine_cls, comm, shard):
        eng = engine_cls(make_model(), comm)
        opt = NGD(eng.optimizer_params(), lr=0.05, momentum=0.9, weight_decay=1e-4)
        tr = Trainer(eng, opt, fwd_loss, clip_norm=10.0)
        per = 16 // world if shard else 16
        for s in range(4):
            lo = rank * per if shard else 0
            tr.step((X[s, lo:lo + per], Y[s, lo:lo + per]))
        if hasattr(eng, "full_state_dict"):
            sd = eng.full_state_dict()
        return torch.cat([v.flatten() for _, v in sorted(sd.items())])

    class Solo(Communicator):           # a communicator that behaves like a single process
        def __init__(self):
            self.group, self.symm, self.world, self.rank = None, None, 1, 0

    ref = run(DataParallelEngine, Solo(), False)
    ddp = run(DataParallelEngine, Communicator(), True)
    fsdp = run(FullyShardedEngine, Communicator(), True)
    # fp32 NGD trajectories amplify summation-order differences (two half-batches vs one batch) to ~1e-3;
    # the two engines themselves must agree tightly
    ok = (torch.allclose(ddp, ref, rtol=1e-2, atol=2e-3) and torch.allclose(fsdp, ref, rtol=1e-2, atol=2e-3)
          and torch.allclose(ddp, fsdp, rtol=1e-4, atol=1e-6))
    if not ok:
        print(f"[rank {rank}] ddp err {(ddp - ref).abs().max():.3e} fsdp err {(fsdp - ref).abs().max():.3e}", flush=True)
    t = torch.tensor([1.0 if ok else 0.0])
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    pdist.cleanup()
    if t.item() != 1.0:
        sys.exit(1)
    if rank == 0:
        print("GLOO WORKER OK", flush=True)


if __name__ == "__main__":
    main()
