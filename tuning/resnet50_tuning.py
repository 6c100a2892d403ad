#!/usr/bin/env python
"""Hyper-parameter sweep driver: ResNet-50 on every 10th CIFAR-10 sample, SGD/StepLR or NGD,
extra flags ``--weight_decay`` and ``--gamma`` (reference: tuning/resnet50_tuning.py)."""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
from fdt.cli import main_resnet  # noqa: E402

if __name__ == "__main__":
    main_resnet(tuning=True)
