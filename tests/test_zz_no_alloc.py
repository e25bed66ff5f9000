"""ddp_sampler_indices_static — the entry point for hosts that cannot allocate device memory.

Who needs it: a feeder thread in a process that has already given all device memory to its own fixed pool, a
plain-C host that links libddp_step.so without an allocator, or a loop that cannot tolerate the device-wide
synchronisation cudaMalloc/cudaFree imply.  They cannot supply the permutation buffer ddp_sampler_indices asks
for, so this variant keeps it in an array inside the library.  The check here: called through the C-ABI twice in
a row (with different arguments, so stale contents of the shared array would show), its output must equal the
regular entry point's bit for bit — the tolerance every integer test in this suite uses.
This file sorts last on purpose: a run that stops at the first failure has already run everything else.
"""
import ctypes

import pytest
import torch

from pytorch_distributed_examples_b200 import _native

pytestmark = pytest.mark.gpu


def _regular(lib, n, world, rank, shuffle, seed, epoch, drop_last, stream):
    ns = lib.ddp_sampler_num_samples(n, world, drop_last)
    perm = torch.empty(n, dtype=torch.int32, device="cuda")
    out = torch.full((ns,), -1, dtype=torch.int64, device="cuda")
    rc = lib.ddp_sampler_indices(n, world, rank, shuffle, seed, epoch, drop_last, perm.data_ptr(), out.data_ptr(), stream)
    assert rc == 0, lib.ddp_last_error()
    return out


def test_static_scratch_variant_matches_regular_entry_point_twice_in_a_row():
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    lib = _native.load_library()
    assert lib.ddp_set_device(0) == 0
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # two back-to-back calls; the second uses a smaller dataset and another epoch, so anything left in the
    # shared array by the first call would corrupt it if the kernel relied on a clean buffer
    calls = [(60000, 8, 5, 1, 0, 3, 0), (10000, 3, 2, 1, 42, 7, 1)]
    outs = []
    for n, world, rank, shuffle, seed, epoch, drop_last in calls:
        ns = lib.ddp_sampler_num_samples(n, world, drop_last)
        out = torch.full((ns,), -1, dtype=torch.int64, device="cuda")     # output buffer: caller-owned, as always
        rc = lib.ddp_sampler_indices_static(n, world, rank, shuffle, seed, epoch, drop_last, out.data_ptr(), stream)
        assert rc == 0, lib.ddp_last_error()
        outs.append(out)
    torch.cuda.synchronize()
    for out, args in zip(outs, calls):
        ref = _regular(lib, *args, stream)
        torch.cuda.synchronize()
        assert torch.equal(out, ref), args
    # unshuffled path and the size limit
    out = torch.empty(5, dtype=torch.int64, device="cuda")
    assert lib.ddp_sampler_indices_static(10, 2, 1, 0, 0, 0, 0, out.data_ptr(), stream) == 0
    torch.cuda.synchronize()
    assert out.tolist() == [1, 3, 5, 7, 9]
    rc = lib.ddp_sampler_indices_static(65537, 1, 0, 1, 0, 0, 0, out.data_ptr(), stream)
    assert rc == -1000 and b"exceeds the built-in scratch" in lib.ddp_last_error()


def test_static_scratch_variant_against_the_committed_golden_digest():
    """An anchor that does not go through the regular entry point: the shard a seeded torch DistributedSampler produced
    when the fixtures were generated (tests/golden/sampler.pt, oracle/gen_golden.py), by sha256 and by its first and
    last 16 indices.  n = 60,000, 8 ranks, rank 5, epoch 3 -- the reference's own dataset size (mnist_horovod.py:38-42)."""
    import hashlib
    import os
    fx = torch.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sampler.pt"), weights_only=False)
    case = next(c for c in fx["cases"] if c["n"] == 60000 and c["world"] == 8 and c["rank"] == 5)
    torch.cuda.set_device(0)
    lib = _native.load_library()
    assert lib.ddp_set_device(0) == 0
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    out = torch.full((case["len"],), -1, dtype=torch.int64, device="cuda")
    rc = lib.ddp_sampler_indices_static(case["n"], case["world"], case["rank"], 1, case["seed"], case["epoch"],
                                        int(case["drop_last"]), out.data_ptr(), stream)
    assert rc == 0, lib.ddp_last_error()
    torch.cuda.synchronize()
    host = out.cpu()
    assert torch.equal(host[:16], case["head"]) and torch.equal(host[-16:], case["tail"])
    assert hashlib.sha256(host.numpy().tobytes()).hexdigest() == case["sha256"]
