"""numpy restatement of the synthetic 28x28 dataset (what replaces the MNIST download of
ref: horovod/mnist_horovod.py:34-38; torchvision ToTensor = u8/255, Normalize = (x-0.1307)/0.3081).

Sample i (class visible in the image, like MNIST, so training on it is well behaved):
    label  = Philox(counter=(i_lo, i_hi, 49, 0), key=seed)[0] % 10
    noise  = little-endian bytes of Philox4x32-10(counter=(i_lo, i_hi, blk, 0), key=seed), blk = 0..48 (16 bytes each)
    stroke = bit k of Philox(counter=(label, blk, 0x7E3D, 0x5EED), key=seed)[0]   (a fixed 50%-dense mask per class)
    pixel  = (noise >> 1) + (112 if stroke else 0)
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import numpy as np

_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = 0x9E3779B9, 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0: int, k1: int):
    """Vectorised over counter arrays (uint64 holding 32-bit values)."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) & _MASK for c in (c0, c1, c2, c3))
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)) & _MASK, lo1, (hi0 ^ c3 ^ np.uint64(k1)) & _MASK, lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return c0, c1, c2, c3


def synth_pixels(seed: int, indices):
    """(uint8 [B,784], int64 [B]) — the raw pixels and labels."""
    idx = np.asarray(indices, dtype=np.uint64)
    b = idx.shape[0]
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    lo, hi = idx & _MASK, idx >> np.uint64(32)
    lab = philox4x32_10(lo, hi, np.full(b, 49, np.uint64), np.zeros(b, np.uint64), k0, k1)[0] % np.uint64(10)
    blk = np.arange(49, dtype=np.uint64)
    c0, c1, c2 = np.repeat(lo, 49), np.repeat(hi, 49), np.tile(blk, b)
    r = philox4x32_10(c0, c1, c2, np.zeros_like(c0), k0, k1)
    words = np.stack(r, axis=1).reshape(b, 49, 4).astype(np.uint32)
    noise = np.stack([(words >> (8 * k)) & 0xFF for k in range(4)], axis=-1).reshape(b, 49, 16)   # little-endian bytes
    st = philox4x32_10(np.repeat(lab, 49), c2, np.full(b * 49, 0x7E3D, np.uint64), np.full(b * 49, 0x5EED, np.uint64),
                       k0, k1)[0].reshape(b, 49).astype(np.uint32)
    stroke = np.stack([(st >> k) & 1 for k in range(16)], axis=-1)                                  # [B,49,16]
    px = (noise >> 1) + np.where(stroke == 1, 112, 0)
    return px.reshape(b, 784).astype(np.uint8), lab.astype(np.int64)


def synth_batch(seed: int, indices):
    """(float32 [B,1,28,28] after ToTensor+Normalize, int64 [B])."""
    px, y = synth_pixels(seed, indices)
    x = (px.astype(np.float32) / np.float32(255.0) - np.float32(0.1307)) / np.float32(0.3081)
    return x.reshape(-1, 1, 28, 28).astype(np.float32), y
