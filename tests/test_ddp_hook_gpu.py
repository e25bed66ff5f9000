"""Row A7 on ONE GPU: stock DistributedDataParallel (ref: pytorch_elastic/mnist_ddp_elastic.py:58) with world_size = 1
fires its communication hooks exactly as it does on 8 GPUs, so the hook plumbing -- CUDA-aware Future, comm stream,
bucket views, optimizer-in-hook with DDP's bucket rebuild after the first iteration -- is exercised by the driver's
single-GPU `pytest -m gpu` run.  (The exchange itself, W > 1, is tests/multi_gpu_worker.py.)
    * pkg.Net  + allreduce_hook       == hook-less DDP  and  == the reference loop's golden parameters
    * pkg.Model + allreduce_hook      == hook-less DDP  and  == the oracle MLP loop
    * torch MLP + FusedUpdateHook     == hook-less DDP + torch.optim.Adam / SGD-momentum
"""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.nn.functional as F

from helpers import RTOL, assert_close_scaled, load_golden
from oracle import mlp_oracle, net_oracle, synth_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def world1(native):
    """A one-rank NCCL process group for DDP + the package's own communicator at size 1."""
    import pytorch_distributed_examples_b200 as pkg
    created = False
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29561")
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
        created = True
    pkg.hvd.init()
    yield pkg
    if created:
        dist.destroy_process_group()


def _ddp(module):
    from torch.nn.parallel import DistributedDataParallel as DDP
    return DDP(module, device_ids=[0], gradient_as_bucket_view=True)


def test_net_under_ddp_with_allreduce_hook_matches_hookless_and_golden(world1):
    pkg = world1
    fx = load_golden("train_sgd.pt")

    def run(hook):
        m = pkg.Net(); m.load_state_dict(fx["state_dict0"]); m = m.cuda().train()
        d = _ddp(m)
        st = None
        if hook:
            st = pkg.HookState()
            d.register_comm_hook(st, pkg.allreduce_hook)
        o = net_oracle.make_optimizer(fx["opt_kind"], d.parameters(), **fx["opt_kw"])
        losses = []
        for k in range(fx["xs"].shape[0]):            # the reference loop body, ref: mnist_ddp_elastic.py:70-75
            o.zero_grad()
            loss = F.nll_loss(d(fx["xs"][k].cuda(), fx["masks2d"][k].cuda(), fx["masks1d"][k].cuda()), fx["ys"][k].cuda())
            loss.backward()
            o.step()
            losses.append(loss.item())
        torch.cuda.synchronize()
        return m, losses, st

    base, l_base, _ = run(False)
    hooked, l_hook, st = run(True)
    assert st.calls >= fx["xs"].shape[0], "the hook must fire once per backward"
    assert l_base == l_hook
    assert torch.equal(base.flat_params, hooked.flat_params), "allreduce_hook at W=1 must leave DDP's result untouched"
    for k, ref in enumerate(fx["losses"]):
        assert abs(l_hook[k] - ref.item()) <= RTOL * abs(ref.item())
    for key, v in hooked.state_dict().items():
        assert_close_scaled(v, fx["state_dict"][key], name=key)


def test_native_mlp_under_ddp_with_hook_matches_hookless_and_oracle(world1):
    pkg = world1
    Hn, Fn, Bn, steps, lr = 2, 128, 32, 3, 0.05
    torch.manual_seed(41)
    sd0 = {k: v.detach().clone() for k, v in mlp_oracle.OracleMLP(Hn, Fn).state_dict().items()}
    xn, yn = synth_oracle.synth_batch(1234, np.arange(steps * Bn))
    xs = torch.from_numpy(xn).view(steps, Bn, 1, 28, 28); ys = torch.from_numpy(yn).view(steps, Bn)
    _, sd_ref = mlp_oracle.train_steps(sd0, xs, ys, Hn, Fn, "sgd", world=1, lr=lr)
    crit = torch.nn.CrossEntropyLoss()

    def run(hook):
        m = pkg.Model(hidden_layers=Hn, features=Fn); m.load_state_dict(sd0); m = m.cuda()
        d = _ddp(m)
        if hook:
            d.register_comm_hook(pkg.HookState(), pkg.allreduce_hook)
        o = torch.optim.SGD(d.parameters(), lr=lr)
        for k in range(steps):
            o.zero_grad()
            crit(d(xs[k].cuda()), ys[k].cuda()).backward()
            o.step()
        torch.cuda.synchronize()
        return m

    base, hooked = run(False), run(True)
    assert torch.equal(base.flat_params, hooked.flat_params)
    g0 = mlp_oracle.forward_backward(sd0, xs[0], ys[0], Hn, Fn, dtype=torch.float64)[2]
    for key, v in hooked.state_dict().items():
        ref = sd_ref[key]
        # SGD maps gradient error linearly into parameters: 1e-5 of the parameter scale + lr*steps*(1e-5 + 4 ReLU flips / B) of the gradient scale
        tol = RTOL * ref.abs().max().item() + lr * steps * (RTOL + 4.0 / Bn) * g0[key].abs().max().item()
        assert (v.cpu() - ref).abs().max().item() <= tol, key


@pytest.mark.parametrize("opt_kind", ["adam", "sgdm"])
def test_fused_update_hook_matches_ddp_plus_torch_optimizer(world1, opt_kind):
    pkg = world1

    def run(fused):
        torch.manual_seed(8)
        m = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(784, 96), torch.nn.ReLU(), torch.nn.Linear(96, 96),
                                torch.nn.ReLU(), torch.nn.Linear(96, 10)).cuda()
        d = _ddp(m)
        o = torch.optim.Adam(d.parameters(), lr=1e-3) if opt_kind == "adam" else \
            torch.optim.SGD(d.parameters(), lr=0.05, momentum=0.9)
        st = None
        if fused:
            st = pkg.FusedUpdateHook.State(o)
            d.register_comm_hook(st, pkg.FusedUpdateHook.hook)
        ds = pkg.SyntheticMNIST(length=4096, seed=4)
        crit = torch.nn.CrossEntropyLoss()
        for k in range(5):                             # > 2: DDP rebuilds its buckets after the first iteration
            x, y = ds.batch(base_index=k * 32, count=32)
            o.zero_grad()
            crit(d(x), y).backward()
            if not fused:
                o.step()
        torch.cuda.synchronize()
        return torch.cat([p.detach().reshape(-1) for p in m.parameters()]).clone(), st

    base, _ = run(False)
    fused, st = run(True)
    assert st.iteration == 5
    scale = base.abs().max().item()
    # same gradients bit-for-bit (same kernels), so only the update rule's arithmetic differs: 1e-5 of the scale
    # (+ 1e-2 of one Adam step for elements whose |g| ~ eps)
    extra = 1e-2 * 1e-3 if opt_kind == "adam" else 0.0
    assert (fused - base).abs().max().item() <= RTOL * scale + extra


def test_buffer_broadcast_rehomes_batchnorm_buffers_without_changing_the_model(world1):
    """N4: BatchNorm buffers live in one flat arena per dtype and are synced by a forward pre-hook; at W = 1 the sync is a
    no-op, so a conv+BN model under DDP(broadcast_buffers=False) + BufferBroadcast must train exactly like stock DDP."""
    pkg = world1

    def run(ours):
        torch.manual_seed(12)
        m = torch.nn.Sequential(torch.nn.Conv2d(1, 8, 3), torch.nn.BatchNorm2d(8), torch.nn.ReLU(), torch.nn.Conv2d(8, 8, 3),
                                torch.nn.BatchNorm2d(8), torch.nn.ReLU(), torch.nn.AdaptiveAvgPool2d(1), torch.nn.Flatten(),
                                torch.nn.Linear(8, 10)).cuda()
        keys = list(m.state_dict().keys())
        bb = pkg.BufferBroadcast(m) if ours else None
        assert list(m.state_dict().keys()) == keys
        from torch.nn.parallel import DistributedDataParallel as DDP
        d = DDP(m, device_ids=[0], broadcast_buffers=not ours)
        o = torch.optim.SGD(d.parameters(), lr=0.05)
        ds = pkg.SyntheticMNIST(length=4096, seed=6)
        for k in range(4):
            x, y = ds.batch(base_index=k * 32, count=32)
            o.zero_grad(); F.cross_entropy(d(x), y).backward(); o.step()
        torch.cuda.synchronize()
        if ours:
            assert bb.calls == 4 and len(bb.arenas) == 2          # float32 statistics + int64 num_batches_tracked
            assert all(b.untyped_storage().data_ptr() in {a.untyped_storage().data_ptr() for a in bb.arenas} for b in m.buffers())
        return {k: v.detach().clone() for k, v in m.state_dict().items()}

    base, mine = run(False), run(True)
    for k in base:            # cuDNN's convolution backward is not bit-reproducible run to run: compare tightly, not bitwise
        if base[k].is_floating_point():
            torch.testing.assert_close(mine[k], base[k], rtol=1e-4, atol=1e-6, msg=k)
        else:
            assert torch.equal(base[k], mine[k]), k
