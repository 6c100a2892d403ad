"""Reference arm of ``bench.py``: the UNMODIFIED reference modules from ``baseline/_ref``
(``resnet.py``, ``transformer.py``, ``ngd_optimizer.py``) driven through their own public API
(``resnet50()``, ``Transformer()``, ``NGD``) inside the reference's own training-loop recipe.

The reference's entry scripts cannot be imported: they train at import time and need torchtext,
madgrad, prefetch_generator, fairscale, ``torch._six`` ....  So the loop bodies of
``resnet50_test.py:521-567`` and ``transformer_test.py:241-283`` are restated here around the
unmodified modules: ``autocast(fp16)`` + ``GradScaler`` + ``clip_grad_norm_(10)`` + ``NGD`` +
(manifold) mixup + the per-step ``loss.item()`` / progress-line host reads, with
``DistributedDataParallel(model, device_ids=[rank])`` (ResNet, ``resnet50_test.py:716``) or
``FullyShardedDataParallel(... CPUOffload(offload_params=True))`` (Transformer,
``transformer_test.py:387-392``) when more than one rank runs.  Nothing from this repository's
models, kernels or engine is on this path.
"""
from __future__ import annotations

import os
import sys
import time
import warnings

import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")


def _import_ref():
    if REF not in sys.path:
        sys.path.insert(0, REF)
    warnings.filterwarnings("ignore")
    import ngd_optimizer  # noqa: F401  (reference module)
    import resnet  # noqa: F401
    import transformer  # noqa: F401
    return resnet, transformer, ngd_optimizer


def _mixup_data(x, y, alpha, device):
    # restated from resnet50_test.py:355-376 (Beta on the host, .item(), device randperm)
    lam = torch.distributions.beta.Beta(alpha, alpha).sample().item() if alpha > 0 else alpha
    index = torch.randperm(x.size(0), device=device)
    return lam * x + (1 - lam) * x[index, :], y, y[index], lam


def _mixup_criterion(criterion, pred, y_a, y_b, lam):
    return lam * criterion(pred, y_a) + (1 - lam) * criterion(pred, y_b)


def _install_fair_mlp(transformer):
    """ref-fair arm: the reference's ``MLPScratch`` computes its ReLU mask with a Python ``for`` over every element of the
    hidden activation (``transformer.py:323-324``: B x 1024 iterations, each a device round trip) — 98 % of its step time.
    This swaps in a subclass whose backward does the SAME arithmetic (including the batch-mean bias gradients) with that
    one loop written as a tensor select.  Nothing else of the reference changes, so a ratio against this arm is not won by
    deleting that loop."""
    base = transformer.MLPScratch

    class MLPScratchVectorised(base):
        @staticmethod
        @torch.amp.custom_bwd(device_type="cuda")          # same autocast state in backward as the reference's decorator gives
        def backward(ctx, g):
            X, W1, b1, W2, b2, pre, post = ctx.saved_tensors
            db2 = g.mean(0, keepdim=True) if b2 is not None else None
            dW2 = g.transpose(-1, -2) @ post
            dpost = g @ W2
            dpre = torch.where(pre > 0, dpost, torch.zeros_like(dpost))       # <- the vectorised loop
            db1 = dpre.mean(0, keepdim=True) if b1 is not None else None
            dW1 = dpre.transpose(-1, -2) @ X
            return dpre @ W1, dW1, db1, dW2, db2

    transformer.MLPScratch = MLPScratchVectorised
    return base


def bench(args, rank, local, world, timed_steps, barrier_sync, max_over_ranks, ClockSampler, fair=False):
    resnet, transformer, ngd_optimizer = _import_ref()
    restore = _install_fair_mlp(transformer) if (fair and args.model == "transformer") else None
    try:
        return _bench(args, rank, local, world, timed_steps, barrier_sync, max_over_ranks, ClockSampler,
                      resnet, transformer, ngd_optimizer)
    finally:
        if restore is not None:
            transformer.MLPScratch = restore


def _bench(args, rank, local, world, timed_steps, barrier_sync, max_over_ranks, ClockSampler, resnet, transformer, ngd_optimizer):
    from torch import autocast
    from torch.amp import GradScaler
    device = torch.device("cuda", local)
    torch.manual_seed(123456)
    torch.backends.cudnn.enabled = True
    torch.backends.cudnn.benchmark = True
    bs = args.bs
    g = torch.Generator().manual_seed(rank)
    criterion = nn.CrossEntropyLoss()
    scaler = GradScaler("cuda")
    state = {"train_loss": torch.zeros(1, device=device), "correct": torch.zeros(1, device=device),
             "total": torch.zeros(1, device=device), "batch_idx": 0, "last": 0.0}

    if args.model == "resnet50":
        net = resnet.resnet50(10).to(device)
        if world > 1:
            from torch.nn.parallel import DistributedDataParallel as DDP
            net = DDP(net, device_ids=[local], output_device=local)
        else:
            net = nn.DataParallel(net, device_ids=[local])
        lr = 0.01 * (world if world > 1 else 1)
        optimizer = ngd_optimizer.NGD(net.parameters(), lr=lr, momentum=0.9, weight_decay=1e-4)
        host = [(torch.randn(bs, 3, 32, 32, generator=g).pin_memory(),
                 torch.randint(0, 10, (bs,), generator=g).pin_memory()) for _ in range(8)]
        dev = [(a.to(device), b.to(device)) for a, b in host]
        h2d = host[0][0].numel() * 4 + host[0][1].numel() * 8

        def body(inputs, targets):
            inputs, targets_a, targets_b, lam = _mixup_data(inputs, targets, 0.99, device)
            optimizer.zero_grad(set_to_none=True)
            with autocast(device_type="cuda", dtype=torch.float16):
                outputs = net(inputs)
                loss = _mixup_criterion(criterion, outputs, targets_a, targets_b, lam)
            scaler.scale(loss).backward()
            scaler.unscale_(optimizer)
            torch.nn.utils.clip_grad_norm_(net.parameters(), 10.)
            scaler.step(optimizer)
            scaler.update()
            state["train_loss"] += loss.item()
            _, predicted = outputs.max(1)
            state["total"] += targets.size(0)
            state["correct"] += (lam * predicted.eq(targets_a.data).sum().float()
                                 + (1 - lam) * predicted.eq(targets_b.data).sum().float())
            state["last"] = "batch idx: {}, Loss: {:.3f} | Acc: {:.3f} ({}/{})".format(
                state["batch_idx"], state["train_loss"].cpu().item() / (state["batch_idx"] + 1),
                100. * (state["correct"] / state["total"]).cpu().item(), int(state["correct"].cpu().item()),
                int(state["total"].cpu().item()))
            state["batch_idx"] += 1
            return loss
    else:
        transformer.device = "cuda"
        model = transformer.Transformer(4, 30522, alpha=0.99).to(device)
        if world > 1:
            from torch.distributed.fsdp import CPUOffload, FullyShardedDataParallel
            from torch.distributed.fsdp.wrap import size_based_auto_wrap_policy
            model = FullyShardedDataParallel(model, auto_wrap_policy=size_based_auto_wrap_policy,
                                             device_id=torch.cuda.current_device(),
                                             cpu_offload=CPUOffload(offload_params=True))
        else:
            model = nn.DataParallel(model, device_ids=[local])
        lr = 1e-4 * (world if world > 1 else 1)
        optimizer = ngd_optimizer.NGD(model.parameters(), lr=lr, weight_decay=0., momentum=0.9)
        L = args.seq_len
        pos_index = torch.arange(start=0, end=512, device=device)
        host = []
        for _ in range(8):
            tok = torch.randint(1000, 30522, (bs, L), generator=g)
            lens = torch.randint(L // 4, L + 1, (bs,), generator=g)
            mask = (torch.arange(L).unsqueeze(0) < lens.unsqueeze(1)).long()
            host.append(((tok * mask).pin_memory(), torch.randint(1, 5, (bs,), generator=g).pin_memory(),
                         torch.zeros(bs, L, dtype=torch.long).pin_memory(), mask.pin_memory()))
        dev = [tuple(t.to(device) for t in b) for b in host]
        h2d = sum(t.numel() * t.element_size() for t in host[0])

        def body(tokens, labels, token_types, masks):
            labels = labels - 1
            optimizer.zero_grad(set_to_none=True)
            with autocast(device_type="cuda", dtype=torch.float16):
                logits, index, lam = model(tokens, token_types, pos_index, masks.view(masks.shape[0], 1, 1, masks.shape[1]))
                loss = _mixup_criterion(criterion, logits, labels, labels[index], lam)
            scaler.scale(loss).backward()
            scaler.unscale_(optimizer)
            torch.nn.utils.clip_grad_norm_(model.parameters(), 10.)
            scaler.step(optimizer)
            scaler.update()
            _, predicted = logits.max(1)
            state["total"] += labels.size(0)
            state["correct"] += (lam * predicted.eq(labels.data).sum().float()
                                 + (1 - lam) * predicted.eq(labels[index].data).sum().float())
            state["last"] = "batch idx: {}, Loss: {:.3f} | Acc: {:.3f}".format(
                state["batch_idx"], float(loss) / (state["batch_idx"] + 1), 100. * (state["correct"] / state["total"]).cpu().item())
            state["batch_idx"] += 1
            return loss

    def step_dev(i):
        body(*dev[i % len(dev)])

    # settle: the reference refreshes every Fisher factor on each of its first 10 steps and every
    # 4th step afterwards (ngd_optimizer.py:347-354); both arms are measured in that steady state
    for i in range(0 if args.no_settle else 12):
        step_dev(i)
    sampler = ClockSampler(local)
    if rank == 0 and not getattr(args, "no_clocks", False):
        sampler.start()
    ms, per = timed_steps(step_dev, args.steps, max(args.warmup, 3), world)

    last = [0.0]

    def step_e2e(i):
        hb = host[i % len(host)]
        db = tuple(t.to(device, non_blocking=True) for t in hb)
        last[0] = body(*db).item()

    for i in range(2):
        step_e2e(i)
    barrier_sync(world)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    s.record()
    for i in range(args.steps):
        step_e2e(i)
    e.record()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    barrier_sync(world)
    e2e_ms = max_over_ranks(max(s.elapsed_time(e) / args.steps, wall_ms), world)
    clocks = sampler.stop() if rank == 0 else {}
    return ms, per, e2e_ms, h2d, 4, 0, clocks, last[0], "fp16 autocast (reference recipe)"
