"""Synthetic datasets (there is no network on the GPU box, so MNIST/CIFAR/ImageNet files cannot be
downloaded the way ref:main.py:107,113 does).  Shapes and dtypes match what the reference's
``ToTensor()+Normalize`` pipeline yields: float32 images, int64 labels."""
from __future__ import annotations

import torch


class SyntheticImages(torch.utils.data.Dataset):
    def __init__(self, n: int, shape=(1, 28, 28), num_classes: int = 10, seed: int = 1234, device="cpu"):
        g = torch.Generator().manual_seed(seed)
        self.labels = torch.randint(0, num_classes, (n,), generator=g)
        # class-dependent mean so a model can actually learn something from it
        base = torch.randn((num_classes,) + tuple(shape), generator=g) * 0.5
        self.images = (base[self.labels] + torch.randn((n,) + tuple(shape), generator=g)).float()
        if device != "cpu":
            self.images, self.labels = self.images.to(device), self.labels.to(device)

    def __len__(self) -> int:
        return self.labels.shape[0]

    def __getitem__(self, i):
        return self.images[i], self.labels[i]


class SyntheticMNIST(SyntheticImages):
    """``[1, 28, 28]`` float32 images, labels in 0..9 (MNIST shape, ref:main.py:107-108)."""

    def __init__(self, n: int = 60000, seed: int = 1234, device="cpu"):
        super().__init__(n, (1, 28, 28), 10, seed, device)


class SyntheticTokens(torch.utils.data.Dataset):
    def __init__(self, n: int, seq_len: int = 1024, vocab: int = 50257, seed: int = 1234):
        g = torch.Generator().manual_seed(seed)
        self.tokens = torch.randint(0, vocab, (n, seq_len + 1), generator=g)

    def __len__(self) -> int:
        return self.tokens.shape[0]

    def __getitem__(self, i):
        return self.tokens[i, :-1], self.tokens[i, 1:]
