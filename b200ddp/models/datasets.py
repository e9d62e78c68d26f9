"""Synthetic data for the extension configs (token sequences, ImageNet-shaped images).

Same contract as ``SyntheticDataset`` (reference ``train.py:53-67``): resident
``data``/``labels`` tensors, generated per process, unseeded unless asked."""
from __future__ import annotations

from typing import Optional

import torch
from torch.utils.data import Dataset


class SyntheticTokens(Dataset):
    """`data[i]` = seq_len token ids, `labels[i]` = the next-token targets (shifted by one)."""

    def __init__(self, num_samples: int, seq_len: int = 1024, vocab_size: int = 50257,
                 seed: Optional[int] = None, device=None):
        gen = torch.Generator().manual_seed(seed) if seed is not None else None
        toks = torch.randint(0, vocab_size, (num_samples, seq_len + 1), generator=gen)
        self.data = toks[:, :-1].contiguous()
        self.labels = toks[:, 1:].contiguous()
        self.num_samples = num_samples
        if device is not None and torch.device(device).type != "cpu":
            self.data, self.labels = self.data.to(device), self.labels.to(device)

    def __len__(self):
        return self.num_samples

    def __getitem__(self, i):
        return self.data[i], self.labels[i]


class SyntheticImages(Dataset):
    def __init__(self, num_samples: int, size: int = 224, num_classes: int = 1000,
                 seed: Optional[int] = None, device=None):
        gen = torch.Generator().manual_seed(seed) if seed is not None else None
        self.data = torch.randn(num_samples, 3, size, size, generator=gen)
        self.labels = torch.randint(0, num_classes, (num_samples,), generator=gen)
        self.num_samples = num_samples
        if device is not None and torch.device(device).type != "cpu":
            self.data, self.labels = self.data.to(device), self.labels.to(device)

    def __len__(self):
        return self.num_samples

    def __getitem__(self, i):
        return self.data[i], self.labels[i]
