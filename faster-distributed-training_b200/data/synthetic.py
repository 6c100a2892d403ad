"""Synthetic CIFAR-10-shape and AG_NEWS-shape datasets (no network on the GPU box).

Shapes and dtypes follow what the reference's pipelines hand to the model: uint8 32×32×3 images
with labels 0..9 (``resnet50_test.py:161-292``), and ``(input_ids, label 1..4, token_type_ids,
attention_mask)`` batches with a BERT-size vocabulary of 30522 (``transformer_test.py:93-104,
141-143``) at a fixed sequence length (BASELINE.json config 3: L = 128).
"""
from __future__ import annotations

import torch


class SyntheticCIFAR:
    """Random images with a learnable signal: class ``c`` brightens a class-specific patch, so loss
    curves are meaningful in end-to-end tests."""

    def __init__(self, n: int = 50000, num_classes: int = 10, seed: int = 0, hw: int = 32):
        g = torch.Generator().manual_seed(seed)
        self.labels = torch.randint(0, num_classes, (n,), generator=g)
        self.images = torch.randint(0, 200, (n, hw, hw, 3), generator=g, dtype=torch.uint8)   # NHWC uint8
        cell = max(hw // 4, 1)
        for c in range(num_classes):
            idx = (self.labels == c).nonzero().squeeze(1)
            r, col = divmod(c, 4)
            r %= 4
            self.images[idx, r * cell:(r + 1) * cell, col * cell:(col + 1) * cell, :] = 255
        self.num_classes = num_classes

    def __len__(self):
        return self.labels.numel()

    def tensors(self):
        return self.images, self.labels


class SyntheticAGNews:
    def __init__(self, n: int = 120000, seq_len: int = 128, vocab: int = 30522, num_classes: int = 4,
                 seed: int = 0, variable_length: bool = True):
        g = torch.Generator().manual_seed(seed)
        self.labels = torch.randint(1, num_classes + 1, (n,), generator=g)            # 1..4 like AG_NEWS
        self.tokens = torch.randint(1000, vocab, (n, seq_len), generator=g, dtype=torch.int32)
        # class signal: a class-specific token repeated at a few positions
        for c in range(1, num_classes + 1):
            idx = (self.labels == c).nonzero().squeeze(1)
            self.tokens[idx, 1:6] = 2000 + c
        self.tokens[:, 0] = 101                                                       # [CLS]
        if variable_length:
            lens = torch.randint(seq_len // 4, seq_len + 1, (n,), generator=g)
        else:
            lens = torch.full((n,), seq_len)
        ar = torch.arange(seq_len).unsqueeze(0)
        self.mask = (ar < lens.unsqueeze(1)).to(torch.int32)
        self.tokens = self.tokens * self.mask                                         # [PAD] = 0
        self.types = torch.zeros(n, seq_len, dtype=torch.int32)
        self.vocab, self.seq_len, self.num_classes = vocab, seq_len, num_classes

    def __len__(self):
        return self.labels.numel()

    def tensors(self):
        return self.tokens, self.labels, self.types, self.mask
