#!/usr/bin/env python
"""ResNet-50 / CIFAR-10 training entry point (same flags as the reference's script of this name).

    python resnet50_test.py --bs 64 --ngd --meta_learning
    python -m torch.distributed.run --nproc_per_node=8 resnet50_test.py --bs 128 --distributed --ngd

All logic lives in ``fdt.cli``; nothing runs at import time.
"""
from fdt.cli import main_resnet

if __name__ == "__main__":
    main_resnet()
