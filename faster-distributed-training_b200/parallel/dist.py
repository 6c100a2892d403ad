"""Process-group bootstrap.

Capability parity with ``utils.py:10-51`` of the reference — explicit-rank set-up, torchrun
``env://`` set-up (the one its drivers use), shared-file rendezvous for multi-node, clean-up and
an ``mp.spawn`` launcher — with the fixes listed in SURVEY.md §8: the device is chosen from
``LOCAL_RANK`` (Q6: the reference uses the global rank), the master address defaults to
``127.0.0.1`` (container hostnames may not resolve), and the backend follows the device
(``nccl`` on GPUs, ``gloo`` on CPU — the reference is NCCL-only and has no CPU path, Q23).
"""
from __future__ import annotations

import datetime
import os
from typing import Callable, Optional

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

DEFAULT_PORT = "12355"  # utils.py:15


def backend_for(device: str) -> str:
    return "nccl" if str(device).startswith("cuda") else "gloo"


def local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", "0")))


def pick_device(prefer_cuda: bool = True) -> torch.device:
    if prefer_cuda and torch.cuda.is_available():
        dev = torch.device("cuda", local_rank() % torch.cuda.device_count())
        torch.cuda.set_device(dev)
        return dev
    return torch.device("cpu")


def setup(rank: int = 0, world_size: int = 1, device: str = "cuda", port: str = DEFAULT_PORT) -> None:
    """Explicit-rank rendezvous (reference ``setup``, utils.py:13-17)."""
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", port)
    dist.init_process_group(backend_for(device), rank=rank, world_size=world_size,
                            timeout=datetime.timedelta(minutes=10))


def setup_norank(world_size: Optional[int] = None, device: str = "cuda") -> None:
    """``env://`` rendezvous under torchrun (reference ``setup_norank``, utils.py:20-23)."""
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", DEFAULT_PORT)
    kw = {}
    if str(device).startswith("cuda") and torch.cuda.is_available():
        kw["device_id"] = torch.device("cuda", local_rank() % torch.cuda.device_count())
    dist.init_process_group(backend_for(device), timeout=datetime.timedelta(minutes=10), **kw)


def setup_sharedfile(world_size: int, rank: int, path: str = "/mnt/nfs/sharedfile", device: str = "cuda") -> None:
    """Shared-file rendezvous for multi-node runs (reference ``setup_sharedfile``, utils.py:26-30)."""
    dist.init_process_group(backend_for(device), init_method=f"file://{path}", world_size=world_size, rank=rank)


def cleanup() -> None:
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()


def is_dist() -> bool:
    return dist.is_available() and dist.is_initialized()


def world_size() -> int:
    return dist.get_world_size() if is_dist() else 1


def rank() -> int:
    return dist.get_rank() if is_dist() else 0


def _spawn_entry(rank_: int, fn: Callable, world: int, device: str, port: str, args):
    os.environ["RANK"] = str(rank_)
    os.environ["LOCAL_RANK"] = str(rank_)
    os.environ["WORLD_SIZE"] = str(world)
    setup(rank_, world, device, port)
    try:
        fn(rank_, world, *args)
    finally:
        cleanup()


def spawn(fn: Callable, world: int, *args, device: str = "cuda", port: str = DEFAULT_PORT) -> None:
    """``mp.spawn`` launcher — the reference's ``distributed_warpper_runner`` (utils.py:37-51),
    generalised: ``fn(rank, world_size, *args)`` runs in every process with the group set up."""
    mp.spawn(_spawn_entry, args=(fn, world, device, port, args), nprocs=world, join=True)
