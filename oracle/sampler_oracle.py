"""Integer restatement of DistributedSampler's index generation (bit-exact contract).

Follows torch: utils/data/distributed.py:102-134 as called by
ref: horovod/mnist_horovod.py:41-42, horovod_mnist_elastic.py:50-51, pytorch_elastic/mnist_ddp_elastic.py:188,89:
    g = Generator(); g.manual_seed(seed + epoch); indices = randperm(len(dataset), generator=g)
    not drop_last: pad with the head of the list (repeating it if needed) up to ceil(N/W)*W
    drop_last:     truncate to floor-ish total_size
    return indices[rank : total_size : W]
and ATen's CPU randperm for n < 2**32/20 (TensorFactories.cpp randperm_cpu):
    r = arange(n); for i in range(n-1): z = mt19937() % (n - i); swap r[i], r[i+z]
with at::mt19937 the standard MT19937 seeded by init_genrand(seed & 0xffffffff).
Pure Python/numpy; validated against torch.randperm / DistributedSampler in tests/test_oracle_cpu.py.
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import math

import numpy as np

_N, _M = 624, 397


class MT19937:
    def __init__(self, seed: int):
        st = np.zeros(_N, dtype=np.uint64)
        s = seed & 0xFFFFFFFF
        st[0] = s
        for j in range(1, _N):
            s = (1812433253 * (s ^ (s >> 30)) + j) & 0xFFFFFFFF
            st[j] = s
        self.state = st.astype(np.uint32)
        self.pos = _N  # forces a twist on first use (at::mt19937 starts with left_=1)

    def _twist(self):
        st = self.state.astype(np.uint64)
        for i in range(_N):
            y = (st[i] & 0x80000000) | (st[(i + 1) % _N] & 0x7FFFFFFF)
            v = st[(i + _M) % _N] ^ (y >> np.uint64(1))
            if int(y) & 1:
                v ^= np.uint64(0x9908B0DF)
            st[i] = v
        self.state = st.astype(np.uint32)
        self.pos = 0

    def block(self) -> np.ndarray:
        """Next 624 tempered outputs (vectorised tempering)."""
        self._twist()
        y = self.state.astype(np.uint64)
        y ^= y >> np.uint64(11)
        y ^= (y << np.uint64(7)) & np.uint64(0x9D2C5680)
        y ^= (y << np.uint64(15)) & np.uint64(0xEFC60000)
        y ^= y >> np.uint64(18)
        return (y & np.uint64(0xFFFFFFFF)).astype(np.uint64)


def randperm(n: int, seed: int) -> np.ndarray:
    r = list(range(n))
    gen = MT19937(seed)
    buf = gen.block()
    k = 0
    for i in range(n - 1):
        if k == _N:
            buf = gen.block()
            k = 0
        z = int(buf[k]) % (n - i)
        k += 1
        r[i], r[i + z] = r[i + z], r[i]
    return np.asarray(r, dtype=np.int64)


def num_samples(n: int, world: int, drop_last: bool) -> int:
    if drop_last and n % world != 0:
        return math.ceil((n - world) / world)
    return math.ceil(n / world)


def shard_indices(n: int, world: int, rank: int, shuffle: bool = True, seed: int = 0, epoch: int = 0,
                  drop_last: bool = False) -> np.ndarray:
    idx = randperm(n, seed + epoch) if shuffle else np.arange(n, dtype=np.int64)
    ns = num_samples(n, world, drop_last)
    total = ns * world
    if not drop_last:
        pad = total - n
        if pad > 0:
            reps = math.ceil(pad / n)
            idx = np.concatenate([idx, np.tile(idx, reps)[:pad]])
    else:
        idx = idx[:total]
    return idx[rank:total:world]
