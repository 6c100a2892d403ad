#!/usr/bin/env python
"""Transformer / AG_NEWS training entry point (same flags as the reference's script of this name).

    python transformer_test.py --batch_size 64 --ngd
    python -m torch.distributed.run --nproc_per_node=8 transformer_test.py --batch_size 64 --distributed --ngd

``--distributed`` selects the fully-sharded engine by default (``--mode ddp`` for data parallel).
"""
from fdt.cli import main_transformer

if __name__ == "__main__":
    main_transformer()
