"""Bit-exact integer/byte paths: shard indices, synthetic data, flatten/unflatten."""
import hashlib

import numpy as np
import pytest
import torch

from helpers import load_golden
from oracle import sampler_oracle, synth_oracle

pytestmark = pytest.mark.gpu


def _indices(native, n, w, r, shuffle, seed, epoch, dl):
    ns = native.sampler_num_samples(n, w, dl)
    perm = torch.empty(max(n, 1), dtype=torch.int32, device="cuda")
    out = torch.empty(ns, dtype=torch.int64, device="cuda")
    native.sampler_indices(n, w, r, shuffle, seed, epoch, dl, perm, out)
    return out.cpu()


def test_sampler_matches_reference_golden(native):
    fx = load_golden("sampler.pt")
    for c in fx["cases"]:
        idx = _indices(native, c["n"], c["world"], c["rank"], True, c["seed"], c["epoch"], c["drop_last"])
        assert idx.numel() == c["len"], c
        assert hashlib.sha256(idx.numpy().tobytes()).hexdigest() == c["sha256"], c


@pytest.mark.parametrize("n,w,seed,epoch,dl", [(60000, 8, 0, 0, False), (60000, 7, 0, 3, False), (10000, 4, 1, 2, True),
                                               (624, 1, 0, 0, False), (625, 2, 3, 0, False), (1, 1, 0, 0, False),
                                               (2, 2, 0, 1, False), (5, 8, 0, 0, False), (65536, 3, 9, 9, False),
                                               (70001, 4, 2, 1, False)])
def test_sampler_matches_torch_distributed_sampler(native, n, w, seed, epoch, dl):
    from torch.utils.data.distributed import DistributedSampler
    for r in sorted({0, w // 2, w - 1}):
        s = DistributedSampler(range(n), num_replicas=w, rank=r, seed=seed, drop_last=dl)
        s.set_epoch(epoch)
        ref = torch.tensor(list(s), dtype=torch.int64)
        assert torch.equal(_indices(native, n, w, r, True, seed, epoch, dl), ref), (n, w, r)
    s = DistributedSampler(range(n), num_replicas=w, rank=w - 1, shuffle=False, drop_last=dl)
    assert torch.equal(_indices(native, n, w, w - 1, False, 0, 0, dl), torch.tensor(list(s), dtype=torch.int64))


def test_sampler_class_and_resharding(native):
    """8 -> 4 -> 8 resharding: every world size partitions the same permutation (ref: elastic get_dataset())."""
    import pytorch_distributed_examples_b200 as pkg
    ds = pkg.SyntheticMNIST(length=60000)
    for w in (8, 4, 8):
        shards = []
        for r in range(w):
            s = pkg.DistributedSampler(ds, num_replicas=w, rank=r)
            s.set_epoch(2)
            assert len(s) == 60000 // w
            shards.append(torch.tensor(list(iter(s))))
        allidx = torch.stack(shards, 1).reshape(-1)          # interleave back: indices[rank::W]
        assert np.array_equal(allidx.numpy(), sampler_oracle.randperm(60000, 2))
    with pytest.raises(ValueError):
        pkg.DistributedSampler(ds, num_replicas=2, rank=2)


def test_synth_batch_bit_exact(native):
    idx = torch.tensor([0, 1, 2, 59999, 123456789012, 7], dtype=torch.int64)
    x = torch.empty(len(idx), 1, 28, 28, device="cuda"); y = torch.empty(len(idx), dtype=torch.int64, device="cuda")
    native.synth_batch(1234, idx.cuda(), 0, len(idx), x, y)
    xo, yo = synth_oracle.synth_batch(1234, idx.numpy())
    assert np.array_equal(x.cpu().numpy(), xo) and np.array_equal(y.cpu().numpy(), yo)
    B = 4096
    x = torch.empty(B, 1, 28, 28, device="cuda"); y = torch.empty(B, dtype=torch.int64, device="cuda")
    native.synth_batch((7 << 32) | 5, None, 1000, B, x, y)
    xo, yo = synth_oracle.synth_batch((7 << 32) | 5, np.arange(1000, 1000 + B))
    assert np.array_equal(x.cpu().numpy(), xo) and np.array_equal(y.cpu().numpy(), yo)
    # golden: the reference fixtures were drawn from this stream
    fx = load_golden("fwdbwd_b8.pt")
    x8 = torch.empty(8, 1, 28, 28, device="cuda"); y8 = torch.empty(8, dtype=torch.int64, device="cuda")
    native.synth_batch(fx["synth_seed"], None, fx["synth_start"], 8, x8, y8)
    assert torch.equal(x8.cpu(), fx["x"]) and torch.equal(y8.cpu(), fx["y"])


def test_u8_pixels_and_device_transform_bit_exact(native):
    """datasets.MNIST holds uint8; ToTensor+Normalize (ref: horovod/mnist_horovod.py:34-37) run on the device."""
    import pytorch_distributed_examples_b200 as pkg
    ds = pkg.SyntheticMNIST(length=60000, seed=1234)
    px, y = ds.batch_u8(base_index=100, count=257)
    x, y2 = ds.batch(base_index=100, count=257)
    assert torch.equal(y, y2)
    xt = pkg.SyntheticMNIST.transform(px)
    assert torch.equal(xt, x), "u8 -> transform must reproduce the float generator bit for bit"
    ref = px.cpu().float().div(255).sub_(0.1307).div_(0.3081).view(-1, 1, 28, 28)      # torchvision's op order
    assert torch.equal(xt.cpu(), ref)
    odd = torch.arange(0, 256, dtype=torch.uint8, device="cuda").repeat(3)[:-5].contiguous()  # n % 16 != 0
    out = torch.empty(odd.numel(), device="cuda")
    native.normalize_u8(odd, out)
    assert torch.equal(out.cpu(), odd.cpu().float().div(255).sub_(0.1307).div_(0.3081))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_bucket_flatten_unflatten_roundtrip(native, dtype):
    from pytorch_distributed_examples_b200.bucket import BucketTable, FlatLayout, flatten_into, unflatten_from
    torch.manual_seed(0)
    shapes = [(10, 1, 5, 5), (10,), (1,), (20, 10, 5, 5), (3, 7), (1024, 1024), (50, 320), (5,), (100003,)]
    base = torch.randn(sum(int(np.prod(s)) for s in shapes) + 64, device="cuda").to(dtype)
    tensors, off = [], 1                       # deliberately odd offsets: exercises the unaligned paths
    for s in shapes:
        n = int(np.prod(s))
        tensors.append(base[off:off + n].view(s)); off += n + 1
    layout = FlatLayout.for_tensors(tensors)
    bucket = torch.zeros(layout.total, dtype=dtype, device="cuda")
    table = BucketTable(torch.device("cuda"))
    flatten_into(bucket, tensors, layout, table)
    for t, v in zip(tensors, layout.views(bucket)):
        assert torch.equal(t, v)
    outs = [torch.zeros_like(t) for t in tensors]
    unflatten_from(bucket, outs, layout, BucketTable(torch.device("cuda")))
    for t, o in zip(tensors, outs):
        assert torch.equal(t, o)
    # empty list is a no-op
    flatten_into(bucket, [], FlatLayout.for_tensors([]), BucketTable(torch.device("cuda")))
