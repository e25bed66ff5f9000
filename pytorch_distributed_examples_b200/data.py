"""Synthetic MNIST-shaped data generated on the device (replaces the torchvision download of
ref: horovod/mnist_horovod.py:34-38 — there is no network; same dtype/shape/normalisation).

``SyntheticMNIST`` is a virtual dataset of ``length`` samples; sample i is a pure function of (seed, i)
(csrc/synth.cu).  ``DeviceLoader`` walks a ``DistributedSampler`` shard in batches and materialises each
batch directly in HBM, yielding ``(data, target)`` like the reference's DataLoader (drop_last=False, so the
last batch may be ragged — e.g. 48 of 128 at world size 2).
"""
from __future__ import annotations

import torch

from . import _native


class SyntheticMNIST(torch.utils.data.Dataset):
    def __init__(self, length=60000, seed=1234, device=None, first_index=0):
        """``first_index``: sample i of this dataset is sample ``first_index + i`` of the virtual stream of ``seed`` --
        a held-out test split is the same seed (same class strokes) with a first_index past the training set."""
        self.length = length
        self.seed = seed
        self.first_index = int(first_index)
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())

    def __len__(self):
        return self.length

    def batch(self, indices: torch.Tensor = None, base_index: int = 0, count: int = None):
        """(x [B,1,28,28] f32, y [B] i64) for dataset indices (int64 CUDA tensor) or base_index..+count."""
        nat = _native.get()
        b = indices.numel() if indices is not None else count
        x = torch.empty(b, 1, 28, 28, dtype=torch.float32, device=self.device)
        y = torch.empty(b, dtype=torch.int64, device=self.device)
        if indices is not None and self.first_index:
            indices = indices + self.first_index
        nat.synth_batch(self.seed, indices, base_index + self.first_index, b, x, y)
        return x, y

    def batch_u8(self, indices: torch.Tensor = None, base_index: int = 0, count: int = None):
        """Raw pixels (uint8 [B,784], int64 labels) — what datasets.MNIST stores before its transform."""
        nat = _native.get()
        b = indices.numel() if indices is not None else count
        px = torch.empty(b, 784, dtype=torch.uint8, device=self.device)
        y = torch.empty(b, dtype=torch.int64, device=self.device)
        if indices is not None and self.first_index:
            indices = indices + self.first_index
        nat.synth_batch_u8(self.seed, indices, base_index + self.first_index, b, px, y)
        return px, y

    @staticmethod
    def transform(pixels_u8: torch.Tensor, mean=0.1307, std=0.3081):
        """ToTensor() + Normalize((mean,), (std,)) on the device: uint8 [B,784] -> float32 [B,1,28,28]."""
        out = torch.empty(pixels_u8.shape[0], 1, 28, 28, dtype=torch.float32, device=pixels_u8.device)
        _native.get().normalize_u8(pixels_u8.contiguous(), out, mean, std)
        return out

    def __getitem__(self, i):
        x, y = self.batch(base_index=int(i), count=1)
        return x[0], y[0]


class DeviceLoader:
    """for data, target in loader — batches born in HBM; mirrors DataLoader(dataset, batch_size, sampler=...)."""

    def __init__(self, dataset: SyntheticMNIST, batch_size: int, sampler=None):
        self.dataset = dataset
        self.batch_size = batch_size
        self.sampler = sampler

    def __len__(self):
        n = len(self.sampler) if self.sampler is not None else len(self.dataset)
        return (n + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        if self.sampler is not None:
            idx = self.sampler.indices_device()
        else:
            idx = torch.arange(len(self.dataset), dtype=torch.int64, device=self.dataset.device)
        for s in range(0, idx.numel(), self.batch_size):
            yield self.dataset.batch(idx[s:s + self.batch_size].contiguous())
