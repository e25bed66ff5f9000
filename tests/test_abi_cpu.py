"""No-GPU checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/ddp_step.h
declares, validates arguments before touching CUDA, and the product never routes through the oracle."""
import ctypes
import os
import re

import pytest
import torch

from helpers import ROOT
from pytorch_distributed_examples_b200 import _native


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "ddp_step.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ddp_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _native.load_library()
    syms = _header_symbols()
    assert len(syms) >= 24
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/ddp_step.h but not exported"
    assert sorted(_native.SIGNATURES) == syms, "ctypes binding table out of sync with the header"
    assert lib.ddp_abi_version() == 1


def test_layout_queries_and_argument_validation_without_gpu():
    nat = _native.Native()
    off, sz = nat.param_layout()
    assert sz == [250, 10, 5000, 20, 16000, 50, 500, 10] and sum(sz) == _native.PARAM_COUNT
    assert off == [0, 252, 264, 5264, 5284, 21284, 21336, 21836] and all(o % 4 == 0 for o in off)
    assert off[-1] + 12 == _native.ARENA_ELEMS
    assert nat.workspace_bytes(64) >= 64 * (1440 * 5 + 320 * 5 + 200) and nat.workspace_bytes(64) % 256 == 0
    assert nat.scratch_bytes(64) > 296 * 21848 * 4
    assert nat.lib.ddp_net_workspace_bytes(0) == -1
    # sampler sizes follow torch (utils/data/distributed.py:88-97)
    from torch.utils.data.distributed import DistributedSampler
    for n, w, dl in [(60000, 8, False), (60000, 7, False), (60000, 7, True), (10, 4, True), (3, 8, True), (3, 8, False)]:
        assert nat.sampler_num_samples(n, w, dl) == DistributedSampler(range(n), w, 0, drop_last=dl).num_samples
    # NULL pointers / bad sizes are rejected with DDP_EINVAL and a message — before any CUDA call
    lib = nat.lib
    rc = lib.ddp_net_forward(None, None, None, None, None, 0, 0, None, 1, 8, None, None, None, None, None, None, None)
    assert rc == -1000 and b"required" in lib.ddp_last_error()
    rc = lib.ddp_net_backward(None, None, None, None, None, 1.0, 1, 8, None, None, None, None)
    assert rc == -1000
    h = _native.Hyper()
    assert lib.ddp_fused_update(0, None, 0, None, None, None, 16, 1.0, ctypes.byref(h), 1, None, None) == -1000
    assert lib.ddp_fused_update(9, 16, 0, 16, None, None, 16, 1.0, ctypes.byref(h), 1, None, None) == -1000
    assert lib.ddp_sampler_indices(0, 1, 0, 1, 0, 0, 0, None, None, None) == -1000
    assert lib.ddp_sampler_indices(10, 2, 2, 1, 0, 0, 0, 16, 16, None) == -1000
    assert lib.ddp_bucket_copy(None, None, None, None, None, 0, None, 0, None) == 0      # empty bucket: no-op
    assert lib.ddp_bucket_copy(None, None, None, None, None, 3, None, 0, None) == -1000
    assert lib.ddp_allreduce_sum(None, None, 4, 0, None) == -1000
    assert lib.ddp_comm_init(None, None, 2, 0) == -1000
    # MLP of the DDP script: layout follows the reference state_dict (ref: pytorch_elastic/mnist_ddp_elastic.py:137-147)
    shape = _native.MLPShape(784, 1024, 5, 10)
    layout, total = nat.mlp_param_layout(shape)
    sizes = [784 * 1024, 1024] + [1024 * 1024, 1024] * 5 + [10 * 1024, 10]      # Model(hidden_layers=5, features=1024)
    assert [n for _, n in layout] == sizes and sum(sizes) == 6_062_090
    assert all(o % 4 == 0 for o, _ in layout) and total == layout[-1][0] + 12
    assert nat.mlp_workspace_bytes(shape, 64) % 256 == 0 and nat.mlp_workspace_bytes(shape, 64) > 64 * 1024 * 4 * 2 * 12
    assert lib.ddp_mlp_workspace_bytes(ctypes.byref(shape), 6) == -1 and b"multiple of 4" in lib.ddp_last_error()
    bad = _native.MLPShape(783, 1024, 5, 10)
    assert lib.ddp_mlp_param_count(ctypes.byref(bad)) == -1
    assert lib.ddp_mlp_forward(ctypes.byref(shape), None, None, None, 64, None, None, None, None) == -1000
    assert lib.ddp_mlp_backward(ctypes.byref(shape), None, None, None, 1.0, 64, None, None, None) == -1000
    assert lib.ddp_split_tf32(None, 4, 4, None, None, None, None, None) == -1000
    assert lib.ddp_gemm_3xtf32(None, None, 0, None, None, 0, 4, 4, 4, None, 4, None) == -1000
    assert lib.ddp_sym_allreduce(None, 0, 0, 2, 0, 0, 4, 0, 1.0, 0, 0, None) == -1000


def test_product_fails_loudly_without_library_or_gpu(tmp_path):
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _native.load_library(os.path.join(tmp_path, "missing.so"))
    import pytorch_distributed_examples_b200 as pkg
    _native.set_instance(None)
    with pytest.raises(RuntimeError, match="CUDA"):
        pkg.Net()(torch.zeros(2, 1, 28, 28))
    with pytest.raises(RuntimeError, match="CUDA"):
        pkg.Model(hidden_layers=1, features=64)(torch.zeros(4, 784))
    nat = _native.get()
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        nat.scale(torch.zeros(4), 2.0)
    with pytest.raises(RuntimeError, match="CUDA"):
        pkg.FusedOptimizer(torch.optim.SGD(torch.nn.Linear(2, 2).parameters(), lr=0.1))


def test_product_never_imports_the_oracle():
    pkg_dir = os.path.join(ROOT, "pytorch_distributed_examples_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f"{f} imports oracle"
                assert "/root/reference" not in text, f"{f} reads the reference tree at run time"
