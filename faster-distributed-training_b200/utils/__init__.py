"""Drop-in names for the reference's ``utils.py``: a user who did
``from utils import *`` finds the same helpers here, backed by the new implementations."""
from ..engine.metrics import draw_graph  # noqa: F401
from ..parallel.ddp import DataParallelEngine as DDP  # noqa: F401
from ..parallel.dist import cleanup, setup, setup_norank, setup_sharedfile, spawn  # noqa: F401
from ..parallel.fsdp import FullyShardedEngine as FSDP  # noqa: F401


def distributed_wrapper(rank, func, world_size, *func_args):
    """Reference ``distributed_wrapper`` (utils.py:37-47): set up the group, wrap the model (the fifth of the
    nine positional arguments) in the data-parallel engine, call ``func``, clean up."""
    assert len(func_args) == 9
    start_epoch, epoch_total, train_dl, test_dl, model, criterion, optimizer, scheduler, alpha = func_args
    import torch
    device = "cuda" if torch.cuda.is_available() else "cpu"
    setup(rank, world_size, device)
    try:
        if device == "cuda":
            torch.cuda.set_device(rank % torch.cuda.device_count())
            model = model.to(torch.device("cuda", rank % torch.cuda.device_count()))
        func(start_epoch, epoch_total, train_dl, test_dl, DDP(model), criterion, optimizer, scheduler, alpha)
    finally:
        cleanup()


def distributed_warpper_runner(distributed_wrapper_fn, world_size, *func_args):
    """Reference spelling kept (utils.py:50-51)."""
    import torch.multiprocessing as mp
    mp.spawn(distributed_wrapper_fn, args=(world_size, *func_args), nprocs=world_size, join=True)
