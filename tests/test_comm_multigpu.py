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


def _torchrun_cli(script, world, port, *flags, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(ROOT / script), "--device", "cuda", "--synthetic", "--no_plots", "--distributed", *flags]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=str(ROOT))
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-3000:])
    return r.stdout


def test_distributed_cli_resnet_meta_mixup_checkpoint_resume(tmp_path):
    """The reference's launch line (run_distributed.sh:2) on 2 GPUs: DDP (sharded step) + NGD + meta-mixup + CUDA graphs,
    sharded evaluation, checkpoint with one optimizer sidecar per rank, then --resume."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    common = ["--bs", "64", "--ngd", "--meta_learning", "--lr", "0.01", "--model", "resnet18", "--steps_per_epoch", "14",
              "--eval_steps", "2", "--cuda_graph", "--ckpt_dir", str(tmp_path)]
    out = _torchrun_cli("resnet50_test.py", 2, 29521, "--epoch", "1", *common)
    assert "test acc" in out
    assert (tmp_path / "resnet_ckpt.pth").exists()
    assert (tmp_path / "resnet_opt_rank0.pth").exists() and (tmp_path / "resnet_opt_rank1.pth").exists()
    out = _torchrun_cli("resnet50_test.py", 2, 29522, "--epoch", "1", "--resume", *common)
    assert "Resuming" in out and "not compatible" not in out


@pytest.mark.parametrize("mode", ["fsdp", "ddp"])
def test_distributed_cli_transformer(tmp_path, mode):
    """run_distributed.sh:3 on 2 GPUs: the transformer under FSDP (the reference's choice) and DDP, variable-length batches."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = _torchrun_cli("transformer_test.py", 2, 29523 if mode == "fsdp" else 29524, "--batch_size", "32", "--ngd", "--epoch", "1",
                        "--steps_per_epoch", "14", "--eval_steps", "2", "--cuda_graph", "--mode", mode, "--ckpt_dir", str(tmp_path))
    assert "test acc" in out
    assert (tmp_path / "transformer_ckpt.pth").exists()
