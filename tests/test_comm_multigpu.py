"""Fused NVLink collectives and the DDP / FSDP engines on >= 2 GPUs (skipped on a single-GPU box).
Each check compares against torch.distributed (NCCL) on the same data."""
import os
import pathlib
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = pathlib.Path(__file__).resolve().parent.parent


def _launch(world, port, env=None, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(ROOT / "tests" / "_comm_worker.py")]
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=str(ROOT), env=e)
    sys.stdout.write(r.stdout[-4000:])
    sys.stderr.write(r.stderr[-3000:])
    assert r.returncode == 0 and "COMM WORKER OK" in r.stdout


@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_collectives_match_nccl(world):
    """Collectives at 1 KB ... 1 GB (incl. an unaligned length), range / bucketed forms, and the engines: symmetric-memory
    path vs NCCL, sharded + overlapped DDP step vs the replicated one."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    _launch(world, 29500 + world)


@pytest.mark.parametrize("env", [{"FDT_FUSED_WGRAD": "1", "FDT_DDP_SHARD": "0"}, {"FDT_AG_PUSH": "1"}],
                         ids=["fused_wgrad", "ag_push"])
def test_optional_comm_paths(env):
    """The opt-in paths (in-switch all-reduce fused into the wgrad epilogue; push all-gather) stay correct."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    _launch(2, 29511, dict(env, COMM_WORKER_MODE="engines" if "FDT_FUSED_WGRAD" in env else "all", COMM_WORKER_SMALL="1"))
