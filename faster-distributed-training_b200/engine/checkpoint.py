"""Checkpoint / resume in the reference's format.

Write: rank 0, only when the test accuracy improves —
``torch.save({'net': state_dict, 'acc': float, 'epoch': int}, './checkpoint/<name>_ckpt.pth')``
(``resnet50_test.py:664-675``, ``transformer_test.py:334-345``), with ``module.``-prefixed keys when
the model is wrapped by an engine (as DDP / DataParallel did in the reference).

Read: ``--resume`` restores ``net``, ``best_acc`` and ``start_epoch``
(``resnet50_test.py:470-475, 680-690``).  Loading accepts prefixed and un-prefixed keys in either
direction (the reference's distributed resume is broken by this mismatch, Q10).

Beyond the reference: an optional sidecar ``<name>_opt.pth`` keeps optimizer / scheduler / RNG
state (the NGD factors are plain tensors here, so they can be saved).  When the engine shards the
optimizer state (FSDP, or DDP with the sharded optimizer step) every rank writes its own
``<name>_opt_rank<r>.pth`` — a rank only holds the factors / momentum of the tensors it owns — and
reads its own file back on ``--resume``.
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Tuple

import torch

CKPT_DIR = "./checkpoint"


def ckpt_path(name: str, directory: str = CKPT_DIR) -> str:
    return os.path.join(directory, f"{name}_ckpt.pth")


def _strip_or_add_prefix(sd: Dict[str, torch.Tensor], want_prefix: bool) -> Dict[str, torch.Tensor]:
    has = all(k.startswith("module.") for k in sd) and len(sd) > 0
    if has == want_prefix:
        return sd
    if want_prefix:
        return {"module." + k: v for k, v in sd.items()}
    return {k[len("module."):]: v for k, v in sd.items()}


def save_checkpoint(name: str, model: torch.nn.Module, acc: float, epoch: int, rank: int = 0,
                    directory: str = CKPT_DIR, optimizer=None, scheduler=None) -> Optional[str]:
    """All ranks must call this (an FSDP model gathers collectively); rank 0 writes."""
    sd_fn = getattr(model, "full_state_dict", model.state_dict)
    sd = {k: v.detach().to("cpu").clone() for k, v in sd_fn().items()}
    sharded = bool(getattr(model, "optimizer_state_sharded", False))
    if optimizer is not None and (rank == 0 or sharded):
        os.makedirs(directory, exist_ok=True)
        side = {"optimizer": optimizer.state_dict(),
                "scheduler": None if scheduler is None else scheduler.state_dict(),
                "rng": torch.get_rng_state(), "rank": int(rank), "sharded": sharded}
        torch.save(side, sidecar_path(name, directory, rank if sharded else None))
    if rank != 0:
        return None
    os.makedirs(directory, exist_ok=True)
    path = ckpt_path(name, directory)
    torch.save({"net": sd, "acc": float(acc), "epoch": int(epoch)}, path)
    return path


def sidecar_path(name: str, directory: str = CKPT_DIR, rank: Optional[int] = None) -> str:
    return os.path.join(directory, f"{name}_opt.pth" if rank is None else f"{name}_opt_rank{rank}.pth")


def load_best_performance(name: str, num_class: int, resume: bool, directory: str = CKPT_DIR) -> Tuple[float, int]:
    best_acc, start_epoch = 1.0 / num_class, 0
    if resume:
        path = ckpt_path(name, directory)
        if not os.path.isfile(path):
            raise FileNotFoundError(f"Error: no checkpoint found at {path}")
        ck = torch.load(path, map_location="cpu", weights_only=False)
        best_acc = max(best_acc, float(ck["acc"]))
        start_epoch = int(ck["epoch"])
    return best_acc, start_epoch


def load_weights(name: str, model: torch.nn.Module, directory: str = CKPT_DIR, strict: bool = True) -> Dict:
    ck = torch.load(ckpt_path(name, directory), map_location="cpu", weights_only=False)
    own = model.state_dict()
    want_prefix = len(own) > 0 and all(k.startswith("module.") for k in own)
    sd = _strip_or_add_prefix(ck["net"], want_prefix)
    with torch.no_grad():
        missing = [k for k in own if k not in sd]
        unexpected = [k for k in sd if k not in own]
        if strict and (missing or unexpected):
            raise KeyError(f"checkpoint mismatch: missing={missing[:5]} unexpected={unexpected[:5]}")
        for k, v in sd.items():
            if k in own:
                own[k].copy_(v.to(own[k].device, own[k].dtype))   # in place: flat-buffer views stay valid
    return ck


def load_optimizer_sidecar(name: str, optimizer, scheduler=None, directory: str = CKPT_DIR, rank: int = 0,
                           sharded: bool = False) -> bool:
    """``sharded``: the engine shards optimizer state — only this rank's own ``_opt_rank<r>.pth`` is acceptable (another
    rank's shard would have the wrong tensors; the replicated ``_opt.pth`` of an unsharded run is tried second and is
    rejected by the optimizer's own validation if it does not fit)."""
    cands = [sidecar_path(name, directory, rank), sidecar_path(name, directory)] if sharded else [sidecar_path(name, directory)]
    path = next((c for c in cands if os.path.isfile(c)), None)
    if path is None:
        return False
    side = torch.load(path, map_location="cpu", weights_only=False)
    if bool(side.get("sharded", False)) and int(side.get("rank", rank)) != int(rank):
        print(f"==> optimizer sidecar {path} belongs to rank {side.get('rank')}; starting with fresh optimizer state")
        return False
    try:
        optimizer.load_state_dict(side["optimizer"])
    except (KeyError, ValueError, AssertionError) as e:
        # the sidecar was written by a different optimizer (e.g. NGD vs MADGRAD): weights resume, state restarts
        print(f"==> optimizer sidecar {path} not compatible with {type(optimizer).__name__} ({type(e).__name__}); "
              f"starting with fresh optimizer state")
        return False
    if scheduler is not None and side.get("scheduler") is not None:
        try:
            scheduler.load_state_dict(side["scheduler"])
        except Exception:
            pass
    return True
