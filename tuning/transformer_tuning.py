#!/usr/bin/env python
"""Hyper-parameter sweep driver: Transformer on every 10th AG_NEWS sample, SGD or NGD with
momentum 0, extra flag ``--weight_decay`` (reference: tuning/transformer_tuning.py)."""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
from fdt.cli import main_transformer  # noqa: E402

if __name__ == "__main__":
    main_transformer(tuning=True)
