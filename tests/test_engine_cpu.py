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


def test_loader_small_ring_and_staged_batches():
    """ring < 8 used to index past the allocated staging buffers (advisor finding): the ring is now as long as it says."""
    import torch
    from fdt.data.loader import TensorBatchLoader
    x = torch.arange(64 * 3).reshape(64, 3).float()
    y = torch.arange(64)
    dl = TensorBatchLoader((x, y), 4, shuffle=False, drop_last=True, pin=False, ring=2)
    assert dl.ring >= 8 and len(dl._stage) == dl.ring
    seen = [b[1].clone() for b in dl]
    assert len(seen) == 16 and torch.equal(torch.cat(seen), y)


def test_optimizer_sidecar_is_per_rank_when_sharded(tmp_path):
    """A sharded engine writes one optimizer sidecar per rank, and a rank never loads another rank's shard."""
    import torch
    from fdt.engine import checkpoint as ckpt
    from fdt.optim import NGD

    class Eng(torch.nn.Module):
        optimizer_state_sharded = True

        def __init__(self):
            super().__init__()
            self.l = torch.nn.Linear(8, 6, bias=False)

    for rank in (0, 1):
        eng = Eng()
        opt = NGD(eng.parameters(), lr=0.1, momentum=0.9)
        eng.l.weight.grad = torch.randn(6, 8)
        opt.step()
        ckpt.save_checkpoint("m", eng, 0.5, 1, rank, str(tmp_path), opt, None)
    assert (tmp_path / "m_opt_rank0.pth").exists() and (tmp_path / "m_opt_rank1.pth").exists()
    assert not (tmp_path / "m_opt.pth").exists()
    eng = Eng()
    opt = NGD(eng.parameters(), lr=0.1, momentum=0.9)
    assert ckpt.load_optimizer_sidecar("m", opt, None, str(tmp_path), rank=1, sharded=True)
    # a shard that belongs to another rank is refused
    import shutil
    shutil.copy(tmp_path / "m_opt_rank0.pth", tmp_path / "m_opt_rank3.pth")
    opt3 = NGD(Eng().parameters(), lr=0.1, momentum=0.9)
    assert not ckpt.load_optimizer_sidecar("m", opt3, None, str(tmp_path), rank=3, sharded=True)


def test_ngd_load_state_dict_validates_before_mutating():
    import torch
    from fdt.optim import NGD
    a = torch.nn.Linear(8, 6, bias=False)
    oa = NGD(a.parameters(), lr=0.1, momentum=0.9)
    a.weight.grad = torch.randn(6, 8)
    oa.step()
    sd = oa.state_dict()
    b = torch.nn.Linear(10, 6, bias=False)           # different shape: the state must be rejected untouched
    ob = NGD(b.parameters(), lr=0.3, momentum=0.5)
    with pytest.raises(ValueError):
        ob.load_state_dict(sd)
    assert ob.param_groups[0]["lr"] == 0.3 and ob.param_groups[0]["momentum"] == 0.5
    c = torch.nn.Linear(8, 6, bias=False)
    oc = NGD(c.parameters(), lr=0.3, momentum=0.5)
    oc.load_state_dict(sd)
    assert oc.param_groups[0]["lr"] == 0.1


def test_fallback_counter_and_strict_mode():
    from fdt.ops import _lib
    n0 = _lib.FALLBACKS.count
    _lib.note_fallback("unit_test_op", "because")
    assert _lib.FALLBACKS.count == n0 + 1 and _lib.FALLBACKS.by_name["unit_test_op: because"] >= 1
    _lib.set_strict(True)
    try:
        with pytest.raises(_lib.NativeCoverageError):
            _lib.note_fallback("unit_test_op", "strict")
    finally:
        _lib.set_strict(False)
        _lib.FALLBACKS.count -= 1
        _lib.FALLBACKS.by_name.pop("unit_test_op: because", None)


def test_c_abi_signatures_cover_every_entry_point():
    """Every FDT_API symbol of the sources has a declaration in include/fdt_b200.h (ctypes argtypes come from there)."""
    import pathlib
    import re
    from fdt.ops import _lib
    root = pathlib.Path(__file__).resolve().parent.parent
    names = set()
    for f in (root / "faster-distributed-training_b200" / "csrc").glob("*.cu"):
        names |= set(re.findall(r"FDT_API int (fdt_\w+)", f.read_text()))
    sigs = _lib.signatures()
    assert names and not (names - set(sigs)), sorted(names - set(sigs))
