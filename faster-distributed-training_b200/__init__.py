"""faster-distributed-training_b200 — B200-native (sm_100a) distributed training framework.

Import it as ``fdt`` (see ``fdt/__init__.py`` at the repository root).
"""
