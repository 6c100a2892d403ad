"""Device-side metrics, event timing and curve output.

The reference syncs the host ≥5 times per step for its tqdm line (``resnet50_test.py:550-566``)
and times epochs with ``time.monotonic()`` (:519,:614).  Here the running loss / λ-weighted
correct / total live in one 3-element device tensor, are reduced across ranks with ONE packed
all-reduce per epoch (N4), and reach the host only when a log line is due.  Steps are timed with
CUDA events and reported as the max over ranks.  Curves go to ``<title>.png`` through matplotlib
when it is installed (``utils.py:54-69``), otherwise to ``<title>.json``.
"""
from __future__ import annotations

import json
import time
from typing import List, Optional, Sequence

import torch


class DeviceMetrics:
    """[loss_sum, correct, total] accumulated without host syncs."""

    def __init__(self, device):
        self.acc = torch.zeros(3, dtype=torch.float32, device=device)
        self.steps = 0
        self.ranks = 1          # how many ranks' loss sums are folded into acc[0] (after reduce)

    def reset(self):
        self.acc.zero_()
        self.steps = 0
        self.ranks = 1

    @torch.no_grad()
    def update(self, loss: torch.Tensor, correct: torch.Tensor, batch: int):
        self.acc[0] += loss.detach().float()
        self.acc[1] += correct.detach().float()
        self.acc[2] += float(batch)
        self.steps += 1

    def reduce(self, comm) -> "DeviceMetrics":
        comm.all_reduce_sum_(self.acc)
        self.ranks = comm.world
        return self

    def read(self):
        """One host sync: returns (mean loss per step, accuracy %, correct, total)."""
        loss_sum, correct, total = self.acc.tolist()
        return loss_sum / max(self.steps * self.ranks, 1), 100.0 * correct / max(total, 1.0), correct, total


class StepTimer:
    """CUDA-event timing (wall-clock fallback on CPU); ``elapsed_ms`` syncs once."""

    def __init__(self, device):
        self.cuda = torch.device(device).type == "cuda"
        self._t0 = None
        self._e0 = self._e1 = None

    def start(self):
        if self.cuda:
            self._e0 = torch.cuda.Event(enable_timing=True)
            self._e1 = torch.cuda.Event(enable_timing=True)
            self._e0.record()
        else:
            self._t0 = time.perf_counter()

    def stop(self):
        if self.cuda:
            self._e1.record()
        else:
            self._t1 = time.perf_counter()

    def elapsed_ms(self) -> float:
        if self.cuda:
            self._e1.synchronize()
            return self._e0.elapsed_time(self._e1)
        return (self._t1 - self._t0) * 1e3

    def elapsed_ms_max_over_ranks(self, comm) -> float:
        ms = self.elapsed_ms()
        if comm is None or comm.world == 1:
            return ms
        dev = "cuda" if self.cuda else "cpu"
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        return float(comm.all_reduce_max_(t)[0])


def peak_memory_gb(device) -> float:
    if torch.device(device).type != "cuda":
        return 0.0
    gb = torch.cuda.max_memory_allocated(device) / 1024 ** 3
    torch.cuda.reset_peak_memory_stats(device)
    return gb


def draw_graph(xs, ys, labels, title: str, metric: str, out_dir: str = ".") -> str:
    """Accuracy / time curves → ``<title>.png`` (matplotlib, Agg) or ``<title>.json``."""
    multi = len(xs) > 0 and isinstance(xs[0], (list, tuple)) or hasattr(xs[0], "__len__")
    series = []
    if multi:
        for x, y, l in zip(xs, ys, labels):
            series.append({"label": str(l), "x": [float(v) for v in x], "y": [float(v) for v in y]})
    else:
        series.append({"label": str(labels), "x": [float(v) for v in xs], "y": [float(v) for v in ys]})
    path = f"{out_dir}/{title}"
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        plt.figure(figsize=(12, 8))
        for s in series:
            plt.plot(s["x"], s["y"], label=s["label"], linewidth=2)
        plt.xticks(series[0]["x"])
        plt.xlabel("Epoch/Iteration")
        plt.ylabel(metric)
        plt.title(title)
        plt.legend()
        plt.grid()
        plt.savefig(path + ".png")
        plt.close()
        return path + ".png"
    except Exception:
        with open(path + ".json", "w") as f:
            json.dump({"title": title, "metric": metric, "series": series}, f)
        return path + ".json"
