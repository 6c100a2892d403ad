"""CPU tier: trainer / checkpoint / sampler / loaders / optimizers / CLI plumbing, and the multi-process
(gloo, world size 2) DDP and FSDP engines."""
import os
import pathlib
import subprocess
import sys

import pytest
import torch

ROOT = pathlib.Path(__file__).resolve().parent.parent


def test_sampler_partitions_and_reshuffles():
    from fdt.parallel.sampler import ShardedSampler
    parts = [ShardedSampler(103, 4, r, shuffle=True, seed=7) for r in range(4)]
    idx = torch.cat([p.indices() for p in parts])
    assert idx.numel() == 104 and set(idx.tolist()) == set(range(103))
    e0 = parts[0].indices().clone()
    parts[0].set_epoch(1)
    assert not torch.equal(e0, parts[0].indices())
    assert len(ShardedSampler(103, 4, 0, drop_last=True)) == 25


def test_tensor_loader_and_background_loader():
    from fdt.data.loader import BackgroundLoader, DevicePrefetcher, TensorBatchLoader
    x, y = torch.arange(100).view(100, 1), torch.arange(100)
    dl = TensorBatchLoader((x, y), 16, shuffle=True, drop_last=True, pin=False)
    seen = []
    for bx, by in DevicePrefetcher(BackgroundLoader(dl), "cpu"):
        assert bx.shape == (16, 1) and torch.equal(bx.view(-1), by)
        seen += by.tolist()
    assert len(seen) == 96 and len(set(seen)) == 96


def test_cifar_pickle_reader_and_gpu_transform(tmp_path):
    import pickle
    import numpy as np
    from fdt.data import cifar
    base = tmp_path / "cifar-10-batches-py"
    base.mkdir()
    rng = np.random.default_rng(0)
    for name in cifar.CIFAR10.train_list + cifar.CIFAR10.test_list:
        with open(base / name, "wb") as f:
            pickle.dump({"data": rng.integers(0, 255, (20, 3072), dtype=np.uint8), "labels": list(range(10)) * 2}, f)
    ds = cifar.CIFAR10(str(tmp_path), train=True)
    assert len(ds) == 100 and ds.images.shape == (100, 32, 32, 3) and ds.images.dtype == torch.uint8
    img, lab = ds[3]
    assert img.shape == (3, 32, 32) and 0 <= img.min() and img.max() <= 1 and lab == 3
    out = cifar.gpu_transform(ds.images[:8], train=False)
    mean, std = torch.tensor(cifar.MEAN).view(1, 3, 1, 1), torch.tensor(cifar.STD).view(1, 3, 1, 1)
    assert torch.allclose(out, (ds.images[:8].permute(0, 3, 1, 2).float() / 255 - mean) / std, atol=1e-6)
    aug = cifar.gpu_transform(ds.images[:8], train=True, generator=torch.Generator().manual_seed(0))
    assert aug.shape == (8, 3, 32, 32) and torch.isfinite(aug).all()


def test_agnews_text_pipeline():
    from fdt.data.agnews import HashWordTokenizer, clean, remove_url, striphtml
    assert striphtml("a <b>bold</b> move") == "a bold move"
    assert "www" not in remove_url("see www.example.com now")
    assert clean("This is <i>the</i> best of the www.news.org stories") == "best stories"
    tok = HashWordTokenizer()
    enc = tok(["wall st. bears claw back", "oil"], padding="longest")
    ids, mask = enc["input_ids"], enc["attention_mask"]
    assert ids.shape == mask.shape and ids[0, 0] == 101 and ids.max() < 30522 and mask[1].sum() == 3
    assert (ids * (1 - mask)).sum() == 0


def test_checkpoint_roundtrip_reference_format(tmp_path):
    from fdt.engine import checkpoint as ck
    from fdt.models.resnet import resnet18
    from fdt.parallel.ddp import DataParallelEngine
    torch.manual_seed(0)
    eng = DataParallelEngine(resnet18(10))
    path = ck.save_checkpoint("resnet", eng, 91.5, 7, 0, str(tmp_path))
    blob = torch.load(path, weights_only=False)
    assert set(blob) == {"net", "acc", "epoch"} and blob["acc"] == 91.5 and blob["epoch"] == 7
    assert all(k.startswith("module.") for k in blob["net"])              # as DDP-wrapped reference checkpoints
    assert blob["net"]["module.conv1.0.conv_weight"].shape == (64, 3, 3, 3)
    best, start = ck.load_best_performance("resnet", 10, True, str(tmp_path))
    assert best == 91.5 and start == 7
    fresh = resnet18(10)                                                  # bare model ← prefixed checkpoint (Q10)
    ck.load_weights("resnet", fresh, str(tmp_path))
    for (k, a), (_, b) in zip(fresh.state_dict().items(), eng.module.state_dict().items()):
        assert torch.equal(a, b), k
    eng2 = DataParallelEngine(resnet18(10))                               # flat-buffer views stay valid after load
    ck.load_weights("resnet", eng2, str(tmp_path))
    assert torch.equal(eng2.flat.params[0].data, eng2.module.conv1[0].conv_weight.data)
    assert eng2.module.conv1[0].conv_weight.data_ptr() >= eng2.flat.param_flat.data_ptr()


def test_trainer_learns_and_madgrad_steps():
    from fdt.cli import ResNetStep
    from fdt.data.synthetic import SyntheticCIFAR
    from fdt.engine.trainer import Trainer
    from fdt.models.resnet import resnet18
    from fdt.optim import MADGRAD, MirrorMADGRAD, NGD
    from fdt.parallel.ddp import DataParallelEngine
    torch.manual_seed(0)
    ds = SyntheticCIFAR(64, seed=0)
    imgs, labels = ds.tensors()
    for make in (lambda p: NGD(p, lr=0.02, momentum=0.9, weight_decay=1e-4), lambda p: MADGRAD(p, lr=1e-3),
                 lambda p: MirrorMADGRAD(p, lr=1e-3)):
        torch.manual_seed(0)
        eng = DataParallelEngine(resnet18(10))
        tr = Trainer(eng, make(eng.optimizer_params()), ResNetStep(0.0, None), clip_norm=10.0)
        losses = [float(tr.step((imgs[:32], labels[:32]))[0]) for _ in range(6)]
        assert losses[-1] < losses[0], losses


def _run_dist(script, world=2, port=29631):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)]
    # FDT_DEBUG_POISON: FSDP turns every non-owned parameter into NaN after each step — the run only matches the
    # single-process one if the all-gather really rewrites all of them
    env = dict(os.environ, OMP_NUM_THREADS="2", CUDA_VISIBLE_DEVICES="", FDT_DEBUG_POISON="1")
    return subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=str(ROOT), env=env)


def test_gloo_ddp_and_fsdp_match_single_process():
    r = _run_dist(ROOT / "tests" / "_gloo_worker.py")
    sys.stdout.write(r.stdout[-2000:])
    sys.stderr.write(r.stderr[-2000:])
    assert r.returncode == 0 and "GLOO WORKER OK" in r.stdout


def test_cli_resnet_cpu_smoke(tmp_path):
    r = subprocess.run([sys.executable, str(ROOT / "resnet50_test.py"), "--device", "cpu", "--synthetic", "--bs", "8", "--ngd",
                        "--meta_learning", "--epoch", "1", "--steps_per_epoch", "2", "--eval_steps", "1", "--model", "resnet18",
                        "--no_plots", "--ckpt_dir", str(tmp_path)], capture_output=True, text=True, timeout=600, cwd=str(ROOT))
    assert r.returncode == 0, r.stderr[-2000:]
    assert (tmp_path / "resnet_ckpt.pth").exists() or "test acc" in r.stdout


def test_fp16_loss_scaler_backs_off_and_grows():
    from fdt.engine.amp import DynamicLossScaler
    sc = DynamicLossScaler("cpu", init_scale=1024.0, growth_interval=3)
    sc.update(torch.tensor(1.0))
    assert sc.get_scale() == 512.0
    for _ in range(3):
        sc.update(torch.tensor(0.0))
    assert sc.get_scale() == 1024.0
    g = torch.ones(4) * 512.0
    sc2 = DynamicLossScaler("cpu", init_scale=512.0)
    sc2.unscale_(g)
    assert torch.allclose(g, torch.ones(4))


def test_resume_with_other_optimizer_skips_sidecar(tmp_path):
    from fdt.engine import checkpoint as ck
    from fdt.models.resnet import resnet18
    from fdt.optim import MADGRAD, NGD
    m = resnet18(10)
    opt = NGD(m.parameters(), lr=0.1, momentum=0.9)
    for p in m.parameters():
        p.grad = torch.zeros_like(p)
    ck.save_checkpoint("resnet", m, 50.0, 1, 0, str(tmp_path), opt, None)
    assert ck.load_optimizer_sidecar("resnet", MADGRAD(m.parameters(), lr=1e-3), None, str(tmp_path)) is False
    assert ck.load_optimizer_sidecar("resnet", NGD(m.parameters(), lr=0.1, momentum=0.9), None, str(tmp_path)) is True
