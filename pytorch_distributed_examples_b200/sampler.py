"""``DistributedSampler`` — drop-in for ``torch.utils.data.distributed.DistributedSampler`` as the reference
uses it (ref: horovod/mnist_horovod.py:41-42, horovod_mnist_elastic.py:50-51,
pytorch_elastic/mnist_ddp_elastic.py:188 + ``set_epoch`` :89), with the shard indices generated on the GPU
by ddp_sampler_indices (csrc/sampler.cu) — bit-identical to torch's MT19937 randperm + pad + ``[rank::W]``.
Same constructor, ``__iter__``, ``__len__``, ``set_epoch``; plus ``indices_device()`` which keeps the shard
on the device for the on-device data generator.
"""
from __future__ import annotations

import os

import torch

from . import _native


def _default_world_rank():
    # the reference passes hvd.size()/hvd.rank() or relies on the process group (torchrun env)
    from . import hvd
    if hvd.is_initialized():
        return hvd.size(), hvd.rank()
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        return torch.distributed.get_world_size(), torch.distributed.get_rank()
    if "WORLD_SIZE" in os.environ and "RANK" in os.environ:
        return int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"])
    raise RuntimeError("Requires distributed package to be available")  # torch's own message


class DistributedSampler(torch.utils.data.Sampler):
    def __init__(self, dataset, num_replicas=None, rank=None, shuffle=True, seed=0, drop_last=False, device=None):
        if num_replicas is None or rank is None:
            w, r = _default_world_rank()
            num_replicas = w if num_replicas is None else num_replicas
            rank = r if rank is None else rank
        if rank >= num_replicas or rank < 0:
            raise ValueError(f"Invalid rank {rank}, rank should be in the interval [0, {num_replicas - 1}]")
        self.dataset = dataset
        self.num_replicas = num_replicas
        self.rank = rank
        self.epoch = 0
        self.drop_last = drop_last
        self.shuffle = shuffle
        self.seed = seed
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        n = len(dataset)
        self.num_samples = _native.get().sampler_num_samples(n, num_replicas, drop_last)
        self.total_size = self.num_samples * self.num_replicas
        self._perm = None

    def indices_device(self) -> torch.Tensor:
        """This rank's shard for the current epoch as an int64 CUDA tensor (no host round trip)."""
        n = len(self.dataset)
        nat = _native.get()
        if self._perm is None or self._perm.numel() < n:
            self._perm = torch.empty(max(n, 1), dtype=torch.int32, device=self.device)
        out = torch.empty(self.num_samples, dtype=torch.int64, device=self.device)
        nat.sampler_indices(n, self.num_replicas, self.rank, self.shuffle, self.seed, self.epoch,
                            self.drop_last, self._perm, out)
        return out

    def __iter__(self):
        return iter(self.indices_device().tolist())

    def __len__(self):
        return self.num_samples

    def set_epoch(self, epoch: int):
        self.epoch = epoch
