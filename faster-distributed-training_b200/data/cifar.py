"""CIFAR-10 dataset and on-device augmentation.

The reference vendors torchvision's ``CIFAR10`` (``resnet50_test.py:161-292``): pickled batches →
uint8 NHWC array → per-sample float CHW tensor → a jit-scripted transform (random crop pad 4,
horizontal flip, normalise — in a randomly permuted order, Q8) run in loader workers.  Here the
dataset stays **uint8 NHWC in host memory**, batches are assembled by one ``index_select``
(:class:`fdt.data.loader.TensorBatchLoader`) and crop → flip → normalise (fixed order) run on
the GPU as part of the fused input kernel, so the H2D traffic is 3 KB per image instead of 12 KB.
Falls back to :class:`SyntheticCIFAR` when the files are absent and cannot be downloaded.
"""
from __future__ import annotations

import os
import pickle
from typing import Optional, Tuple

import numpy as np
import torch

from .download import check_integrity, download_and_extract_archive
from .synthetic import SyntheticCIFAR

MEAN = (0.4914, 0.4822, 0.4465)
STD = (0.2023, 0.1994, 0.2010)
CLASSES = ("plane", "car", "bird", "cat", "deer", "dog", "frog", "horse", "ship", "truck")


def get_classes():
    return CLASSES


class CIFAR10:
    base_folder = "cifar-10-batches-py"
    url = "https://www.cs.toronto.edu/~kriz/cifar-10-python.tar.gz"
    filename = "cifar-10-python.tar.gz"
    tgz_md5 = "c58f30108f718f92721af3b95e74349a"
    train_list = [f"data_batch_{i}" for i in range(1, 6)]
    test_list = ["test_batch"]

    def __init__(self, root: str = "./data", train: bool = True, download: bool = False):
        self.root, self.train = os.path.expanduser(root), train
        if download and not self._present():
            download_and_extract_archive(self.url, self.root, filename=self.filename, md5=self.tgz_md5)
        if not self._present():
            raise RuntimeError("Dataset not found. You can use download=True to download it")
        data, labels = [], []
        for name in (self.train_list if train else self.test_list):
            with open(os.path.join(self.root, self.base_folder, name), "rb") as f:
                entry = pickle.load(f, encoding="latin1")
            data.append(entry["data"])
            labels.extend(entry.get("labels", entry.get("fine_labels")))
        arr = np.vstack(data).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1)      # NHWC uint8
        self.images = torch.from_numpy(np.ascontiguousarray(arr))
        self.labels = torch.tensor(labels, dtype=torch.long)
        self.num_classes = 10

    def _present(self) -> bool:
        return all(check_integrity(os.path.join(self.root, self.base_folder, n))
                   for n in self.train_list + self.test_list)

    def __len__(self):
        return self.labels.numel()

    def tensors(self):
        return self.images, self.labels

    def __getitem__(self, i) -> Tuple[torch.Tensor, int]:
        """Per-sample access with the reference's convention: float CHW in [0, 1]."""
        return self.images[i].permute(2, 0, 1).float() / 255.0, int(self.labels[i])


def get_dataset(root: str = "./data", synthetic: bool = False, n_train: int = 50000, n_test: int = 10000):
    """(train, test) datasets; synthetic stand-ins when requested or when CIFAR-10 is unavailable."""
    if not synthetic:
        try:
            import torch.distributed as dist
            multi = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
            if multi and dist.get_rank() != 0:
                dist.barrier()          # rank 0 downloads / extracts first (the reference lets every rank race, resnet50_test.py:311)
            try:
                out = CIFAR10(root, True, download=True), CIFAR10(root, False, download=True)
            finally:
                if multi and dist.get_rank() == 0:
                    dist.barrier()
            return out
        except Exception as e:  # no files and no network
            print(f"==> CIFAR-10 unavailable ({type(e).__name__}); using synthetic CIFAR-shape data")
    return SyntheticCIFAR(n_train, seed=0), SyntheticCIFAR(n_test, seed=1)


_CONST = {}


def _mean_std(dev):
    """Per-device normalisation constants, created once (no H2D copy inside CUDA-graph capture)."""
    key = str(dev)
    if key not in _CONST:
        _CONST[key] = (torch.tensor(MEAN, device=dev).view(1, 3, 1, 1), torch.tensor(STD, device=dev).view(1, 3, 1, 1))
    return _CONST[key]


def gpu_transform(images_u8: torch.Tensor, train: bool, generator: Optional[torch.Generator] = None,
                  pad: int = 4) -> torch.Tensor:
    """uint8 NHWC → float NCHW, random-crop(pad 4) + horizontal flip (train) + normalise.

    Pure-torch batched definition (CPU path and oracle); the CUDA path fuses the same maths with
    mixup and the bf16/NHWC conversion in one kernel."""
    B, H, W, C = images_u8.shape
    dev = images_u8.device
    x = images_u8.permute(0, 3, 1, 2).float() / 255.0
    mean, std = _mean_std(dev)
    if train:
        dx = torch.randint(0, 2 * pad + 1, (B,), device=dev, generator=generator) - pad
        dy = torch.randint(0, 2 * pad + 1, (B,), device=dev, generator=generator) - pad
        flip = torch.rand(B, device=dev, generator=generator) < 0.5
        ys = torch.arange(H, device=dev).view(1, H, 1) + dy.view(B, 1, 1)
        xs = torch.arange(W, device=dev).view(1, 1, W) + dx.view(B, 1, 1)
        xs = torch.where(flip.view(B, 1, 1), (W - 1) - xs + 2 * dx.view(B, 1, 1), xs)
        valid = (ys >= 0) & (ys < H) & (xs >= 0) & (xs < W)
        ysc, xsc = ys.clamp(0, H - 1), xs.clamp(0, W - 1)
        bidx = torch.arange(B, device=dev).view(B, 1, 1)
        g = x[bidx.expand(B, H, W), :, ysc.expand(B, H, W), xsc.expand(B, H, W)]   # [B,H,W,C]
        g = g * valid.expand(B, H, W).unsqueeze(-1)
        x = g.permute(0, 3, 1, 2)
    return (x - mean) / std
