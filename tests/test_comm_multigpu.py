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


@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_collectives_match_nccl(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    port = 29500 + world
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(ROOT / "tests" / "_comm_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=420, cwd=str(ROOT))
    sys.stdout.write(r.stdout[-3000:])
    sys.stderr.write(r.stderr[-3000:])
    assert r.returncode == 0 and "COMM WORKER OK" in r.stdout
