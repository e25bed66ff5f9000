"""Flat arenas and the flatten-into-bucket table.

Replaces what DDP's C++ Reducer does per bucket (torch: nn/parallel/distributed.py:1255-1280: copy each
``param.grad`` into a slot of one contiguous buffer, allreduce it, copy back) and Horovod's fusion buffer
(ref: horovod/mnist_horovod.py:53).  Layout rule: tensors in the given order, each slot start rounded up
to 4 elements (16 bytes for fp32) so the flatten kernel and the fused update use 128-bit accesses.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence

import torch

from . import _native

ALIGN_ELEMS = 4


@dataclass
class FlatLayout:
    offsets: List[int]     # element offset of each tensor in the arena
    sizes: List[int]       # numel of each tensor
    shapes: List[torch.Size]
    total: int             # arena length in elements (multiple of ALIGN_ELEMS)

    @staticmethod
    def for_tensors(tensors: Sequence[torch.Tensor]) -> "FlatLayout":
        offsets, sizes, shapes, off = [], [], [], 0
        for t in tensors:
            offsets.append(off)
            sizes.append(t.numel())
            shapes.append(t.shape)
            off += (t.numel() + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS
        return FlatLayout(offsets, sizes, shapes, off)

    def views(self, arena: torch.Tensor) -> List[torch.Tensor]:
        return [arena[o:o + n].view(s) for o, n, s in zip(self.offsets, self.sizes, self.shapes)]


def arena_of(tensors: Sequence[torch.Tensor], layout: FlatLayout):
    """If ``tensors`` already alias one contiguous buffer exactly as ``layout`` describes, return
    (base_ptr); else None.  Used to adopt Net's own arenas zero-copy (gradients born in the bucket)."""
    if not tensors:
        return None
    t0 = tensors[0]
    if t0 is None or not t0.is_contiguous():
        return None
    base = t0.data_ptr() - layout.offsets[0] * t0.element_size()
    for t, off in zip(tensors, layout.offsets):
        if t is None or not t.is_contiguous() or t.dtype != t0.dtype or t.device != t0.device:
            return None
        if t.data_ptr() != base + off * t.element_size():
            return None
    return base


class BucketTable:
    """Device-resident (ptr, offset, nbytes, chunk) tables for ddp_bucket_copy — built once per layout and
    reused every step; rebuilt only when a tensor's address changes (e.g. fresh ``.grad`` allocations)."""

    def __init__(self, device):
        self.device = device
        self._key = None
        self.nchunks = 0
        self.ptrs = self.offsets = self.nbytes = self.chunk_item = self.chunk_start = None

    def update(self, tensors: Sequence[torch.Tensor], layout: FlatLayout, elem_size: int):
        key = tuple(t.data_ptr() for t in tensors)
        if key == self._key:
            return self
        ptrs, offs, nb, citem, cstart = [], [], [], [], []
        for i, (t, off, n) in enumerate(zip(tensors, layout.offsets, layout.sizes)):
            if not t.is_contiguous():
                raise RuntimeError("gradient tensors must be contiguous to be flattened")
            ptrs.append(t.data_ptr())
            offs.append(off * elem_size)
            nbytes = n * t.element_size()
            nb.append(nbytes)
            for start in range(0, nbytes, _native.BUCKET_CHUNK_BYTES):
                citem.append(i)
                cstart.append(start)
        dev = self.device
        # uint64 pointers are shipped as int64 bit patterns
        self.ptrs = torch.tensor([p if p < 2 ** 63 else p - 2 ** 64 for p in ptrs], dtype=torch.int64, device=dev)
        self.offsets = torch.tensor(offs, dtype=torch.int64, device=dev)
        self.nbytes = torch.tensor(nb, dtype=torch.int64, device=dev)
        self.chunk_item = torch.tensor(citem, dtype=torch.int32, device=dev)
        self.chunk_start = torch.tensor(cstart, dtype=torch.int64, device=dev)
        self.nchunks = len(citem)
        self._key = key
        self._tensors = list(tensors)                 # host-side mirror (kept alive; used by test doubles)
        self._offsets_elems = list(layout.offsets)
        return self


def flatten_into(bucket: torch.Tensor, tensors: Sequence[torch.Tensor], layout: FlatLayout, table: BucketTable):
    """tensors -> bucket (one kernel launch)."""
    table.update(tensors, layout, bucket.element_size())
    _native.get().bucket_copy(table, bucket, 0)


def unflatten_from(bucket: torch.Tensor, tensors: Sequence[torch.Tensor], layout: FlatLayout, table: BucketTable):
    """bucket -> tensors (one kernel launch)."""
    table.update(tensors, layout, bucket.element_size())
    _native.get().bucket_copy(table, bucket, 1)
