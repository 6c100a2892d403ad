#!/usr/bin/env python
"""Headline benchmark: ResNet-50 (CIFAR-10 shape) training throughput, NGD + mixup, bf16, DDP.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the unmodified reference from baseline/_ref

Metric (BASELINE.json): images / second over the whole job, device-timed (CUDA events), max over
ranks, weak scaling (128 images per GPU; 8 GPUs = the reference's global batch 1024).  The
transformer classifier (AG_NEWS shape, L=128) is measured the same way with ``--model transformer``.

Two numbers per run:
* ``value`` — device-timed training steps (forward, backward, gradient all-reduce, clip, NGD
  optimizer step — nothing skipped) on inputs already resident on the device, each step timed by
  its own event pair with an L2 flush (256 MB write) between steps;
* ``e2e`` — the same steps driven through the public API a user calls (``fdt.engine.Trainer`` fed
  by ``fdt.data.loader``), where every step copies its uint8 batch from pinned host memory to the
  device and the step's loss is read back to the host.

The last line printed by rank 0 is the JSON record the driver parses.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BASELINE_IMG_PER_S = 640.0      # ResNet-50/CIFAR-10 "with tricks" ≈ 78 s/epoch ≈ 640 img/s (digitised figure)
BASELINE_SAMPLES_PER_S = 67.0   # Transformer/AG_NEWS "with tricks" ≈ 67 samples/s


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples nvidia-smi SM clocks and throttle reasons every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self._stop, self._t = gpu_index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.idx)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.splitlines()[0].split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm = [float(r[1]) for r in self.rows if len(r) > 2 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for n, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


# ----------------------------------------------------------------------------- helpers
def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def init_dist(world):
    import torch.distributed as dist
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "12355")
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    return dist


def barrier_sync(world):
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x: float, world) -> float:
    import torch.distributed as dist
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


class L2Flusher:
    def __init__(self, mb=256):
        self.buf = torch.empty(mb * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")

    def __call__(self):
        self.buf.fill_(1.0)


def timed_steps(step_fn, steps, warmup, world, flush=True):
    """W untimed steps, then exactly K steps, each with its own CUDA-event pair; L2 flushed between
    steps (outside the event pairs).  Returns (ms_per_step max over ranks, per-step list)."""
    flusher = L2Flusher() if flush else None
    for i in range(warmup):
        step_fn(i)
    barrier_sync(world)
    evs = []
    for i in range(steps):
        if flusher:
            flusher()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        step_fn(warmup + i)
        e.record()
        evs.append((s, e))
    barrier_sync(world)
    per = [s.elapsed_time(e) for s, e in evs]
    return max_over_ranks(sum(per) / len(per), world), per


# ----------------------------------------------------------------------------- our arm
_HEAP = {}


def _heap(device, world, rank):
    """One NVLink symmetric heap per process, shared by every engine this run builds."""
    if world == 1:
        return None
    if "heap" not in _HEAP:
        try:
            from fdt.parallel.symm import SymmetricHeap
            _HEAP["heap"] = SymmetricHeap.create(device)
        except Exception as e:
            _HEAP["heap"] = None
            if rank == 0:
                print(f"[bench] symmetric memory unavailable ({type(e).__name__}: {e}); NCCL collectives", file=sys.stderr)
    return _HEAP["heap"]


def build_fdt(model_name, bs, device, world, use_graph, mode, meta=False):
    from fdt.cli import ResNetStep, TransformerStep
    from fdt.engine.trainer import Trainer
    from fdt.models.resnet import resnet50
    from fdt.models.transformer import Transformer
    from fdt.optim import NGD
    from fdt.parallel.comm import Communicator
    from fdt.parallel.ddp import DataParallelEngine
    from fdt.parallel.fsdp import FullyShardedEngine
    comm = Communicator()
    allocator = None
    heap = _heap(device, world, comm.rank)
    if heap is not None:
        comm.symm = heap
        allocator = heap.alloc_f32
    torch.manual_seed(123456)
    if model_name == "resnet50":
        model = resnet50(10).to(device)
        mm = None
        if meta:        # --meta_learning: learnable per-sample mixing weights (resnet50_test.py:388-401, 525), fused kernels
            from fdt.data.mixup import MetaMixup
            mm = MetaMixup(bs, compat=False, device=device)
        step_obj = ResNetStep(0.99, mm)
        lr, wd = 0.01 * world, 1e-4
    else:
        model = Transformer(4, 30522, alpha=0.99).to(device)
        step_obj = TransformerStep()
        lr, wd = 1e-4 * world, 0.0
    Engine = FullyShardedEngine if (mode == "fsdp" and world > 1) else DataParallelEngine
    engine = Engine(model, comm, allocator=allocator)
    params = list(engine.optimizer_params())
    if model_name == "resnet50" and meta:
        params.append(step_obj.meta.lam)
    opt = NGD(params, lr=lr, momentum=0.9, weight_decay=wd)
    trainer = Trainer(engine, opt, step_obj, clip_norm=10.0, use_cuda_graph=use_graph)
    return trainer, comm


def make_host_batches(model_name, bs, n, seq_len, seed):
    """Pinned host batches in the dtypes the public loaders ship (uint8 images / int32 tokens)."""
    from fdt.data.loader import TensorBatchLoader
    from fdt.data.synthetic import SyntheticAGNews, SyntheticCIFAR
    if model_name == "resnet50":
        ds = SyntheticCIFAR(bs * n, seed=seed)
    else:
        ds = SyntheticAGNews(bs * n, seq_len, seed=seed, variable_length=True)
    return TensorBatchLoader(ds.tensors(), bs, shuffle=True, drop_last=True, pin=True, ring=max(n, 4), seed=seed)


def measure_fdt(args, rank, local, world, model, mode, bs, steps, sample_clocks, meta=False):
    """Device-timed steps + the end-to-end loop for one (model, parallel mode).  Returns a dict."""
    from fdt import ops
    device = torch.device("cuda", local)
    trainer, comm = build_fdt(model, bs, device, world, args.cuda_graph, mode, meta)
    loader = make_host_batches(model, bs, 8, args.seq_len, seed=rank)
    host_batches = [tuple(b) for b in iter(loader)]
    dev_batches = [tuple(t.to(device) for t in b) for b in host_batches]
    launches = ops.launch_counter()

    def step_dev(i):
        trainer.step(dev_batches[i % len(dev_batches)])

    # warm-up long enough to pass the optimizer's 10 always-refresh steps and capture graphs
    warm = max(args.warmup, 3)
    pre = 0 if args.no_settle else 12
    for i in range(pre):
        step_dev(i)
    sampler = ClockSampler(local)
    if rank == 0 and sample_clocks:
        sampler.start()
    l0 = launches.count
    ms, per = timed_steps(step_dev, steps, warm, world)
    n_launch = (launches.count - l0) * steps // (steps + warm)       # launches of the timed steps only
    # ---- end to end through the public API: pinned H2D every step, loss D2H every step
    loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
    h2d_bytes = sum(t.numel() * t.element_size() for t in host_batches[0])

    def step_e2e(i):
        hb = host_batches[i % len(host_batches)]
        db = tuple(t.to(device, non_blocking=True) for t in hb)
        loss, _ = trainer.step(db)
        loss_host.copy_(loss.detach().reshape(1), non_blocking=True)

    for i in range(3):
        step_e2e(i)
    barrier_sync(world)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    s.record()
    for i in range(steps):
        step_e2e(i)
    e.record()
    torch.cuda.synchronize()
    _ = float(loss_host[0])
    wall_ms = (time.perf_counter() - t0) * 1e3 / steps
    barrier_sync(world)
    e2e_ms = max_over_ranks(max(s.elapsed_time(e) / steps, wall_ms), world)
    clocks = sampler.stop() if (rank == 0 and sample_clocks) else {}
    out = {"ms": ms, "per": per, "e2e_ms": e2e_ms, "h2d": h2d_bytes, "d2h": 4, "launches": n_launch, "clocks": clocks,
           "last_loss": float(loss_host[0]), "graph": bool(trainer.use_graph), "steps": steps,
           "sharded_step": bool(getattr(trainer.engine, "sharded", mode == "fsdp" and world > 1))}
    del trainer, dev_batches
    torch.cuda.empty_cache()
    return out


def _series(m, bs, world, unit):
    gb = bs * world
    return {"value": round(gb / (m["ms"] / 1e3), 2), "unit": unit, "ms_per_step": round(m["ms"], 4), "global_batch": gb,
            "steps": m["steps"],
            "e2e": {"value": round(gb / (m["e2e_ms"] / 1e3), 2), "unit": unit, "ms_per_step": round(m["e2e_ms"], 4),
                    "h2d_bytes_per_step": m["h2d"], "d2h_bytes_per_step": m["d2h"]}}


def fdt_extras(args, rank, local, world):
    """The other series BASELINE.json names (transformer samples/s; FSDP), measured in the same process right after
    the headline so they land in the driver's record: same timing rules, fewer steps."""
    steps = max(3, min(args.steps, 10))
    plan = [("transformer_ddp", "transformer", "ddp", 64, False)]
    if world > 1:
        plan += [("transformer_fsdp", "transformer", "fsdp", 64, False), ("resnet50_fsdp", "resnet50", "fsdp", 128, False),
                 # BASELINE.json config 4: ResNet-50 FSDP FULL_SHARD, NGD + meta-mixup, 256 img/GPU
                 ("resnet50_fsdp_meta_bs256", "resnet50", "fsdp", 256, True)]
    else:
        plan += [("resnet50_meta_mixup", "resnet50", "ddp", 128, True)]
    out = {}
    for key, model, mode, bs, meta in plan:
        try:
            m = measure_fdt(args, rank, local, world, model, mode, bs, steps, False, meta)
            out[key] = _series(m, bs, world, "img/s" if model == "resnet50" else "samples/s")
            out[key]["parallelism"] = f"{mode}{world}"
        except Exception as e:                      # an auxiliary series must never take the headline down
            out[key] = {"error": f"{type(e).__name__}: {e}"[:300]}
    return out


# ----------------------------------------------------------------------------- reference arm
def run_reference(args, rank, local, world, fair=False):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isfile(os.path.join(ref_dir, "ngd_optimizer.py")):
        return None, "baseline/_ref missing (run python baseline/install_ref.py where /root/reference is mounted)"
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import run_ref
    return run_ref.bench(args, rank, local, world, timed_steps, barrier_sync, max_over_ranks, ClockSampler, fair=fair), None


def reference_extras(args, rank, local, world):
    """Transformer on the reference arm, single GPU only (its multi-GPU path is FSDP with CPU offload and NGD on a
    host-resident shard — minutes per step): ``ref_as_is`` and ``ref_fair`` (only the per-element Python loop of
    MLPScratch.backward vectorised, see baseline/run_ref.py)."""
    if world != 1:
        return {}
    import copy
    out = {}
    a = copy.copy(args)
    a.model, a.bs, a.steps, a.warmup, a.no_clocks = "transformer", 64, 3, 3, True
    for key, fair in (("transformer_ref_as_is", False), ("transformer_ref_fair", True)):
        try:
            res, why = run_reference(a, rank, local, world, fair=fair)
            if res is None:
                out[key] = {"unavailable": why}
                continue
            ms, per, e2e_ms = res[0], res[1], res[2]
            out[key] = {"value": round(64 / (ms / 1e3), 2), "unit": "samples/s", "ms_per_step": round(ms, 3), "steps": 3,
                        "e2e": {"value": round(64 / (e2e_ms / 1e3), 2), "unit": "samples/s", "ms_per_step": round(e2e_ms, 3)}}
        except Exception as e:
            out[key] = {"error": f"{type(e).__name__}: {e}"[:300]}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="fdt", choices=["fdt", "reference"])
    ap.add_argument("--model", default="resnet50", choices=["resnet50", "transformer"])
    ap.add_argument("--bs", type=int, default=None, help="per-GPU batch (default 128 resnet50 / 64 transformer)")
    ap.add_argument("--seq_len", type=int, default=128)
    ap.add_argument("--mode", default="ddp", choices=["ddp", "fsdp"])
    ap.add_argument("--cuda_graph", type=int, default=int(os.environ.get("FDT_BENCH_GRAPH", "1")))
    ap.add_argument("--no_settle", action="store_true")
    ap.add_argument("--no_extra", action="store_true", help="skip the auxiliary series (transformer, FSDP) in `extra`")
    args = ap.parse_args()
    if args.bs is None:
        args.bs = 128 if args.model == "resnet50" else 64
    rank, local, world = dist_env()
    if args.impl == "reference" and not os.path.isdir(os.path.join(ROOT, "baseline", "_ref")):
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref not present"}))
        return 0
    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device visible"}))
        return 0
    torch.cuda.set_device(local)
    init_dist(world)
    metric = "resnet50_cifar_train_img_per_s" if args.model == "resnet50" else "transformer_agnews_train_samples_per_s"
    base = BASELINE_IMG_PER_S if args.model == "resnet50" else BASELINE_SAMPLES_PER_S
    if args.impl == "reference":
        res, why = run_reference(args, rank, local, world)
        if res is None:
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": why}))
            return 0
        ms, per, e2e_ms, h2d, d2h, n_launch, clocks, last_loss, dtype = res
        impl_detail = {"cuda_graph": False, "loop": "reference recipe (autocast fp16 + GradScaler + clip + NGD), stock DDP/FSDP"}
        extra = {} if (args.no_extra or args.model != "resnet50") else reference_extras(args, rank, local, world)
    else:
        m = measure_fdt(args, rank, local, world, args.model, args.mode, args.bs, args.steps, True)
        ms, per, e2e_ms, h2d, d2h, n_launch, clocks, last_loss = (m["ms"], m["per"], m["e2e_ms"], m["h2d"], m["d2h"],
                                                                   m["launches"], m["clocks"], m["last_loss"])
        dtype = "bf16"
        impl_detail = {"cuda_graph": m["graph"], "sharded_optimizer_step": m["sharded_step"],
                       "collectives": "in-kernel NVLS/P2P over symmetric memory" if _HEAP.get("heap") is not None else "none (1 GPU)" if world == 1 else "NCCL"}
        extra = {} if (args.no_extra or args.model != "resnet50" or args.mode != "ddp") else fdt_extras(args, rank, local, world)
    gb = args.bs * world
    value = gb / (ms / 1e3)
    rec = {
        "metric": metric, "value": round(value, 2), "unit": "img/s" if args.model == "resnet50" else "samples/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ms, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": round(value / base, 3), "dtype": dtype,
        "data": "synthetic (CIFAR-10-shape uint8 32x32x3 / AG_NEWS-shape tokens; random-init weights)",
        "impl": args.impl,
        "config": {"model": "resnet50-cifar (FusedConvBN variant)" if args.model == "resnet50" else "transformer-6L-d512-h8 classifier",
                   "global_batch": gb, "per_gpu_batch": args.bs, "seq_len": args.seq_len if args.model == "transformer" else None,
                   "parallelism": f"{args.mode}{world}", "optimizer": "NGD(momentum 0.9) + mixup + clip 10",
                   "l2": "256 MB flush write between timed steps; per-step working set >> 126 MB L2"},
        "impl_detail": impl_detail,
        "e2e": {"value": round(gb / (e2e_ms / 1e3), 2), "unit": "img/s" if args.model == "resnet50" else "samples/s",
                "ms_per_step": round(e2e_ms, 4), "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": n_launch, "clocks": clocks, "last_loss": last_loss,
        "step_ms_min_max": [round(min(per), 4), round(max(per), 4)],
        "extra": extra,
    }
    if rank == 0:
        print(json.dumps(rec))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
