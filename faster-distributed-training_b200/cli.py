"""Command-line drivers.

``resnet50_test.py`` / ``transformer_test.py`` / ``tuning/*.py`` at the repository root are thin
entry points into this module.  Flags and defaults are those of the reference
(``resnet50_test.py:46-59``, ``transformer_test.py:350-361``, ``tuning/resnet50_tuning.py:44-59``,
``tuning/transformer_tuning.py:304-316``); the extra flags (``--device``, ``--mode``,
``--synthetic``, ``--cuda-graph`` ...) expose what is new.  Unlike the reference nothing runs at
import time, ``--device cpu`` works (gloo), LR scaling follows the real world size (Q5) and the
device comes from ``LOCAL_RANK`` (Q6).
"""
from __future__ import annotations

import argparse
import os
import sys
import time
from typing import Optional

import numpy as np
import torch

from . import ops
from .data import cifar as cifar_data
from .data import mixup as mx
from .data.agnews import AGNewsDataset, get_tokenizer_size
from .data.loader import DevicePrefetcher, TensorBatchLoader
from .engine import checkpoint as ckpt
from .engine.metrics import DeviceMetrics, StepTimer, draw_graph, peak_memory_gb
from .engine.trainer import Trainer
from .models import resnet as resnet_models
from .models.transformer import Transformer
from .optim import MADGRAD, NGD, MirrorMADGRAD
from .optim.schedules import build_scheduler, scale_lr
from .parallel import dist as pdist
from .parallel.comm import Communicator
from .parallel.ddp import DataParallelEngine
from .parallel.fsdp import FullyShardedEngine
from .parallel.sampler import ShardedSampler


# ----------------------------------------------------------------------------- shared
def _common_flags(p: argparse.ArgumentParser):
    p.add_argument("--device", default="auto", choices=["auto", "cuda", "cpu"])
    p.add_argument("--mode", default=None, choices=["ddp", "fsdp"], help="parallel engine when --distributed")
    p.add_argument("--synthetic", action="store_true", help="CIFAR/AG_NEWS-shape synthetic data")
    p.add_argument("--steps_per_epoch", default=0, type=int, help="cap steps per epoch (0 = full epoch)")
    p.add_argument("--eval_steps", default=0, type=int, help="cap eval batches (0 = full test set)")
    p.add_argument("--cuda_graph", action="store_true", help="capture the train step in CUDA graphs")
    p.add_argument("--dtype", default="bf16", choices=["bf16", "fp16"],
                   help="bf16 (default): no loss scaling needed.  fp16: the reference's AMP flow (resnet50_test.py:533-548) — "
                        "dynamic loss scaling through backward, unscale, found-inf skip, scale update, all on the device; "
                        "the sm_100a kernels keep bf16 storage (fp32 exponent range), so this exercises the GradScaler "
                        "logic rather than half-precision storage")
    p.add_argument("--lr_scale", default=None, type=float, help="LR multiplier under --distributed (reference: 4)")
    p.add_argument("--log_every", default=50, type=int)
    p.add_argument("--seed", default=123456, type=int)
    p.add_argument("--ckpt_dir", default="./checkpoint")
    p.add_argument("--no_plots", action="store_true")


def _resolve_device(args) -> torch.device:
    want_cuda = args.device in ("auto", "cuda") and torch.cuda.is_available()
    if args.device == "cuda" and not torch.cuda.is_available():
        raise SystemExit("--device cuda requested but no CUDA device is visible")
    return pdist.pick_device(want_cuda)


def _wrap(model, args, device, default_mode):
    comm = Communicator()
    allocator = None
    if args.distributed and device.type == "cuda" and comm.world > 1:
        try:
            from .parallel.symm import SymmetricHeap
            heap = SymmetricHeap.create(device)
            comm.symm = heap
            allocator = heap.alloc_f32
        except Exception as e:  # symmetric memory unavailable: library collectives
            if comm.rank == 0:
                print(f"==> symmetric memory unavailable ({type(e).__name__}: {e}); using NCCL collectives")
    mode = args.mode or default_mode
    if args.distributed and mode == "fsdp":
        return FullyShardedEngine(model, comm, allocator=allocator), comm
    return DataParallelEngine(model, comm, allocator=allocator), comm


def _make_scaler(args, device):
    """``--dtype fp16`` → the device-side GradScaler equivalent; bf16 needs none."""
    if getattr(args, "dtype", "bf16") != "fp16":
        return None
    from .engine.amp import DynamicLossScaler
    return DynamicLossScaler(device)


def _progress(tag, epoch, step, nsteps, metrics: DeviceMetrics):
    loss, acc, correct, total = metrics.read()
    print(f"[{tag}] epoch {epoch} batch idx: {step}/{nsteps}, Loss: {loss:.3f} | Acc: {acc:.3f} ({int(correct)}/{int(total)})",
          flush=True)


# ----------------------------------------------------------------------------- ResNet / CIFAR-10
def parse_resnet(argv=None, tuning=False):
    p = argparse.ArgumentParser(description="B200-native CIFAR10 training")
    p.add_argument("--lr", default=0.02, type=float, help="learning rate")
    p.add_argument("--resume", "-r", action="store_true", help="resume from checkpoint")
    p.add_argument("--epoch", default=50, type=int, help="epoch num for training")
    p.add_argument("--alpha", default=0.99, type=float, help="alpha value for beta distribution")
    p.add_argument("--bs", default=128, type=int)
    p.add_argument("--workers", default=2, type=int)
    p.add_argument("--meta_learning", action="store_true", help="adaptive lambda in mixup")
    p.add_argument("--distributed", action="store_true")
    p.add_argument("--ngd", action="store_true")
    p.add_argument("--meta_compat", action="store_true", help="reference behaviour: fresh random lambda every step")
    p.add_argument("--model", default="resnet50", choices=["resnet18", "resnet34", "resnet50", "resnet101", "resnet152"])
    if tuning:
        p.add_argument("--weight_decay", default=5e-4, type=float)
        p.add_argument("--gamma", default=0.75, type=float)
        p.add_argument("--subset", default=10, type=int, help="keep every N-th sample")
    _common_flags(p)
    return p.parse_args(argv)


def get_model(args, num_classes, device):
    print("==> Building model..")
    net = getattr(resnet_models, args.model)(num_classes).to(device)
    return net


def get_optimizer(net_params, args, world, tuning=False):
    lr = scale_lr(args.lr, args.distributed, world, args.lr_scale)
    if tuning:
        wd = args.weight_decay
        if args.ngd:
            opt = NGD(net_params, lr=lr, momentum=0.9, weight_decay=wd)
            return opt, build_scheduler("multistep", opt, milestones=[10, 15], gamma=0.1)
        opt = torch.optim.SGD(net_params, lr=lr, momentum=0.9, weight_decay=wd)
        return opt, build_scheduler("step", opt, step_size=2, gamma=args.gamma)
    if args.ngd:
        opt = NGD(net_params, lr=lr, momentum=0.9, weight_decay=1e-4)
        return opt, build_scheduler("multistep", opt, milestones=[10, 20], gamma=0.2)
    opt = MADGRAD(net_params, lr=lr, momentum=0.9, weight_decay=5e-6)
    return opt, build_scheduler("cosine", opt, T_max=200)


class ResNetStep:
    """forward+loss closure for the Trainer: GPU augmentation → (meta-)mixup → model → fused CE."""

    def __init__(self, alpha, meta: Optional[mx.MetaMixup]):
        self.alpha, self.meta = alpha, meta

    def __call__(self, engine, batch):
        images, targets = batch
        targets = targets.long()
        if self.meta is None and images.dtype == torch.uint8 and ops.use_native(images):
            # one kernel: crop + flip + normalise of both partners, λ-blend, bf16 NHWC out
            lam, index = mx.sample_mixup(self.alpha, images.size(0), images.device)
            x = ops._cu().input_aug_mixup(images, True, index, lam, cifar_data.MEAN, cifar_data.STD)
            ya, yb = targets, targets[index]
            logits = engine(x)
            loss = mx.mixup_criterion(logits, ya, yb, lam)
            with torch.no_grad():
                correct = mx.mixup_correct(logits, ya, yb, lam)
            return loss, correct
        if self.meta is not None and images.dtype == torch.uint8 and ops.use_native(images):
            # meta-mixup: one kernel for augmentation + per-sample sigmoid(theta) blend, dtheta from its backward kernel
            x, ya, yb, lam = self.meta.fused_from_uint8(images, targets, cifar_data.MEAN, cifar_data.STD)
            lam = lam.reshape(-1)
            logits = engine(x)
            loss = mx.mixup_criterion(logits, ya, yb, lam)
            with torch.no_grad():
                correct = mx.mixup_correct(logits, ya, yb, lam)
            return loss, correct
        x = cifar_data.gpu_transform(images, train=True) if images.dtype == torch.uint8 else images
        if self.meta is not None:
            x, ya, yb, lam = self.meta(x, targets)
            lam = lam.reshape(-1)
        else:
            x, ya, yb, lam = mx.mixup_data(x, targets, self.alpha)
        logits = engine(x)
        loss = mx.mixup_criterion(logits, ya, yb, lam)
        with torch.no_grad():
            correct = mx.mixup_correct(logits, ya, yb, lam)
        return loss, correct


@torch.no_grad()
def evaluate_resnet(engine, loader, device, comm, max_steps=0):
    engine.eval()
    acc = torch.zeros(2, device=device)
    for i, (images, targets) in enumerate(DevicePrefetcher(loader, device)):
        if max_steps and i >= max_steps:
            break
        x = cifar_data.gpu_transform(images, train=False)
        logits = engine(x)
        acc[0] += logits.argmax(1).eq(targets.long()).sum()
        acc[1] += targets.numel()
    comm.all_reduce_sum_(acc)
    c, t = acc.tolist()
    engine.train()
    return 100.0 * c / max(t, 1.0)


def run_resnet(args, tuning=False):
    device = _resolve_device(args)
    if args.distributed:
        pdist.setup_norank(device=device.type)
    torch.manual_seed(args.seed)
    rank, world = pdist.rank(), pdist.world_size()
    classes = cifar_data.get_classes()
    best_acc, start_epoch = ckpt.load_best_performance("resnet", len(classes), args.resume, args.ckpt_dir)
    if rank == 0:
        print("==> Preparing data..")
    trainset, testset = cifar_data.get_dataset("./data", args.synthetic)
    tr_t, te_t = trainset.tensors(), testset.tensors()
    if tuning and args.subset > 1:
        tr_t = tuple(t[:: args.subset] for t in tr_t)
        te_t = tuple(t[:: args.subset] for t in te_t)
    pin = device.type == "cuda"
    tr_s = ShardedSampler(tr_t[0].shape[0], world, rank, shuffle=True, seed=args.seed)
    te_s = ShardedSampler(te_t[0].shape[0], world, rank, shuffle=False)      # Q9: eval is sharded
    trainloader = TensorBatchLoader(tr_t, args.bs, tr_s, drop_last=True, pin=pin)
    testloader = TensorBatchLoader(te_t, args.bs, te_s, shuffle=False, drop_last=False, pin=pin)

    net = get_model(args, len(classes), device)
    if args.resume:
        print("==> Resuming from checkpoint..")
        ckpt.load_weights("resnet", net, args.ckpt_dir)
    engine, comm = _wrap(net, args, device, "ddp")
    params = engine.optimizer_params()
    meta = None
    if args.meta_learning:
        meta = mx.MetaMixup(args.bs, compat=args.meta_compat, device=device)
        if not args.meta_compat:
            params = list(params) + [meta.lam]          # λ is really learned (Q3)
    optimizer, scheduler = get_optimizer(params, args, world, tuning)
    if args.resume:
        ckpt.load_optimizer_sidecar("resnet", optimizer, scheduler, args.ckpt_dir, rank,
                                    bool(getattr(engine, "optimizer_state_sharded", False)))
    trainer = Trainer(engine, optimizer, ResNetStep(args.alpha, meta), clip_norm=10.0, use_cuda_graph=args.cuda_graph,
                      scaler=_make_scaler(args, device))

    n_epochs = args.epoch
    training_acc, testing_acc = np.zeros(n_epochs), []
    training_time_epoch = np.zeros(n_epochs)
    metrics = DeviceMetrics(device)
    for epoch in range(start_epoch, start_epoch + n_epochs):
        engine.train()
        trainloader.set_epoch(epoch)                   # Q7
        metrics.reset()
        timer = StepTimer(device)
        timer.start()
        nsteps = len(trainloader) if not args.steps_per_epoch else min(len(trainloader), args.steps_per_epoch)
        for bi, batch in enumerate(DevicePrefetcher(trainloader, device)):
            if bi >= nsteps:
                break
            loss, correct = trainer.step(batch)
            metrics.update(loss, correct, batch[1].numel())
            if rank == 0 and args.log_every and (bi + 1) % args.log_every == 0:
                _progress("train", epoch, bi + 1, nsteps, metrics)
        timer.stop()
        ms = timer.elapsed_ms_max_over_ranks(comm)
        metrics.reduce(comm)
        loss_m, acc_m, _, total = metrics.read()
        training_time_epoch[epoch - start_epoch] = ms / 1e3
        training_acc[epoch - start_epoch] = acc_m
        if rank == 0:
            print(f"epoch {epoch}: train loss {loss_m:.4f} acc {acc_m:.3f}% | {ms/1e3:.2f} s "
                  f"({total / max(ms / 1e3, 1e-9):.0f} img/s, max over ranks)")
            print("Peak memory allocated: {:.2f} GB".format(peak_memory_gb(device)))
        acc = evaluate_resnet(engine, testloader, device, comm, args.eval_steps)
        testing_acc.append(acc)
        if rank == 0:
            print(f"epoch {epoch}: test acc {acc:.3f}%")
        improved = acc > best_acc
        if improved:                                   # all ranks call (collective-safe); rank 0 writes
            if rank == 0:
                print("Saving..")
            ckpt.save_checkpoint("resnet", engine, acc, epoch, rank, args.ckpt_dir, optimizer, scheduler)
            best_acc = acc
        scheduler.step()
    if rank == 0 and not args.no_plots:
        xs = np.arange(start_epoch, start_epoch + n_epochs)
        draw_graph([xs, xs], [training_acc, np.asarray(testing_acc)], ["training", "testing"],
                   "Resnet accuracy curve", "accuracy")
        draw_graph(xs, training_time_epoch, "training time", "Resnet time for training", "time(sec.)")
    if args.distributed:
        pdist.cleanup()
    return {"best_acc": best_acc, "train_acc": training_acc, "test_acc": testing_acc, "time": training_time_epoch}


# ----------------------------------------------------------------------------- Transformer / AG_NEWS
def parse_transformer(argv=None, tuning=False):
    p = argparse.ArgumentParser(description="B200-native AG_NEWS transformer training")
    p.add_argument("--batch_size", "-b", help="batch size", default=128, type=int)
    p.add_argument("--epoch", help="epoch", default=50, type=int)
    p.add_argument("--lr", help="learning rate", default=1e-4, type=float)
    p.add_argument("--resume", help="continue training", action="store_true")
    p.add_argument("--workers", help="number of workers", default=2, type=int)
    p.add_argument("--alpha", help="alpha for beta distribution", default=0.99, type=float)
    p.add_argument("--distributed", help="whether to turn on distributed training", action="store_true")
    p.add_argument("--ngd", action="store_true")
    p.add_argument("--seq_len", default=128, type=int, help="sequence length of the synthetic corpus")
    p.add_argument("--mask_compat", action="store_true", help="reference's -1e-9 mask fill (padding unmasked)")
    p.add_argument("--bias_grad_compat", action="store_true", help="reference's batch-mean head bias gradients")
    p.add_argument("--onecycle_per_step", action="store_true", help="step OneCycleLR every batch instead of every epoch")
    if tuning:
        p.add_argument("--weight_decay", default=5e-4, type=float)
        p.add_argument("--subset", default=10, type=int)
    _common_flags(p)
    return p.parse_args(argv)


class TransformerStep:
    def __call__(self, engine, batch):
        tokens, labels, types, masks = batch
        labels = labels.long() - 1                      # AG_NEWS labels are 1..4 (transformer_test.py:242)
        logits, index, lam = engine(tokens.long(), types.long(), None, masks)
        la, lb = labels, labels[index]
        loss = mx.mixup_criterion(logits, la, lb, lam)
        with torch.no_grad():
            correct = mx.mixup_correct(logits, la, lb, lam)
        return loss, correct


@torch.no_grad()
def evaluate_transformer(engine, loader, device, comm, max_steps=0):
    engine.eval()
    acc = torch.zeros(2, device=device)
    for i, (tokens, labels, types, masks) in enumerate(DevicePrefetcher(loader, device)):
        if max_steps and i >= max_steps:
            break
        logits = engine(tokens.long(), types.long(), None, masks)      # eval returns logits only (Q13)
        acc[0] += logits.argmax(1).eq(labels.long() - 1).sum()
        acc[1] += labels.numel()
    comm.all_reduce_sum_(acc)
    c, t = acc.tolist()
    engine.train()
    return 100.0 * c / max(t, 1.0)


def run_transformer(args, tuning=False):
    device = _resolve_device(args)
    if args.distributed:
        pdist.setup_norank(device=device.type)
    torch.manual_seed(args.seed)
    rank, world = pdist.rank(), pdist.world_size()
    num_class = 4
    vocab = get_tokenizer_size()
    best_acc, start_epoch = ckpt.load_best_performance("transformer", num_class, args.resume, args.ckpt_dir)
    n_train, n_test = (120000, 7600)
    if tuning and args.subset > 1:
        n_train, n_test = n_train // args.subset, n_test // args.subset
    ds = AGNewsDataset(batch_size=args.batch_size, synthetic=args.synthetic, seq_len=args.seq_len,
                       n_train=n_train, n_test=n_test)
    train_dl, test_dl, train_sampler = ds.load_data(args.distributed, args.workers, rank, world, device)

    model = Transformer(num_class, vocab, alpha=args.alpha, mask_compat=args.mask_compat,
                        bias_grad_mean_compat=args.bias_grad_compat).to(device)
    if args.resume:
        print("==> Resuming from checkpoint..")
        ckpt.load_weights("transformer", model, args.ckpt_dir)
    engine, comm = _wrap(model, args, device, "fsdp")   # the reference's distributed transformer path is FSDP
    lr = scale_lr(args.lr, args.distributed, world, args.lr_scale)
    params = engine.optimizer_params()
    wd = args.weight_decay if tuning else 0.0
    if args.ngd:
        optimizer = NGD(params, lr=lr, weight_decay=wd, momentum=0.0 if tuning else 0.9)
    elif tuning:
        optimizer = torch.optim.SGD(params, lr=lr, weight_decay=wd, momentum=0.0)
    else:
        optimizer = MirrorMADGRAD(params, lr=lr, weight_decay=0.0, momentum=0.9)
    epochs_total = args.epoch
    steps_per_epoch = len(train_dl) if not args.steps_per_epoch else min(len(train_dl), args.steps_per_epoch)
    if tuning:
        scheduler = build_scheduler("multistep", optimizer, milestones=[10, 15], gamma=0.1)
    else:
        scheduler = build_scheduler("onecycle", optimizer, max_lr=lr * 5, epochs=max(epochs_total - start_epoch, 1),
                                    steps_per_epoch=max(steps_per_epoch, 1),
                                    cycle_momentum="momentum" in optimizer.defaults and optimizer.defaults["momentum"] > 0)
    if args.resume:
        ckpt.load_optimizer_sidecar("transformer", optimizer, scheduler, args.ckpt_dir, rank,
                                    bool(getattr(engine, "optimizer_state_sharded", False)))
    trainer = Trainer(engine, optimizer, TransformerStep(), clip_norm=10.0, use_cuda_graph=args.cuda_graph,
                      scaler=_make_scaler(args, device))

    training_accuracy, testing_accuracy = [], []
    training_time_epoch = np.zeros(max(epochs_total - start_epoch, 1))
    metrics = DeviceMetrics(device)
    for epoch in range(start_epoch, epochs_total):
        engine.train()
        if train_sampler is not None:
            train_sampler.set_epoch(epoch)
        metrics.reset()
        timer = StepTimer(device)
        timer.start()
        for bi, batch in enumerate(DevicePrefetcher(train_dl, device)):
            if bi >= steps_per_epoch:
                break
            loss, correct = trainer.step(batch)
            metrics.update(loss, correct, batch[1].numel())
            if args.onecycle_per_step and not tuning:
                scheduler.step()
            if rank == 0 and args.log_every and (bi + 1) % args.log_every == 0:
                _progress("train", epoch, bi + 1, steps_per_epoch, metrics)
        timer.stop()
        ms = timer.elapsed_ms_max_over_ranks(comm)
        metrics.reduce(comm)
        loss_m, acc_m, _, total = metrics.read()
        training_time_epoch[epoch - start_epoch] = ms / 1e3
        training_accuracy.append(acc_m)
        if rank == 0:
            print(f"epoch {epoch}: train loss {loss_m:.4f} acc {acc_m:.3f}% | {ms/1e3:.2f} s "
                  f"({total / max(ms / 1e3, 1e-9):.0f} samples/s, max over ranks)")
        if not args.onecycle_per_step or tuning:
            scheduler.step()                            # once per epoch, as the reference does (Q16)
        acc = evaluate_transformer(engine, test_dl, device, comm, args.eval_steps)
        testing_accuracy.append(acc)
        if rank == 0:
            print(f"epoch {epoch}: test acc {acc:.3f}%")
        if acc > best_acc:
            if rank == 0:
                print("Saving..")
            ckpt.save_checkpoint("transformer", engine, acc, epoch, rank, args.ckpt_dir, optimizer, scheduler)
            best_acc = acc
    if rank == 0:
        print("Peak memory allocated: {:.2f} GB".format(peak_memory_gb(device)))
        if not args.no_plots and training_accuracy:
            xs = np.arange(start_epoch, start_epoch + len(training_accuracy))
            draw_graph([xs, xs], [training_accuracy, testing_accuracy], ["training", "testing"],
                       "Transformer accuracy curve", "accuracy")
            draw_graph(xs, training_time_epoch[: len(xs)], "training time", "Transformer time for training", "time(sec.)")
    if args.distributed:
        pdist.cleanup()
    return {"best_acc": best_acc, "train_acc": training_accuracy, "test_acc": testing_accuracy, "time": training_time_epoch}


def main_resnet(argv=None, tuning=False):
    return run_resnet(parse_resnet(argv, tuning), tuning)


def main_transformer(argv=None, tuning=False):
    return run_transformer(parse_transformer(argv, tuning), tuning)
