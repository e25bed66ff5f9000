"""world_size-2 gloo tests (CPU) of the N>1 HOST LOGIC: hvd.init bootstrap through the store, broadcast,
DistributedOptimizer == 'average the gradients then step', elastic commit/restore/sync and an
8->4->8-style resize (here 2->1->2) with LR rescale — the kernels are replaced by tests/fake_native.py."""
import copy
import math
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn
import torch.nn.functional as F


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _setup(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), LOCAL_RANK=str(rank),
                      WORLD_SIZE=str(world), OMP_NUM_THREADS="1")
    torch.set_num_threads(1)
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (here, os.path.dirname(here)):
        if p not in sys.path:
            sys.path.insert(0, p)
    from fake_native import FakeNative
    from pytorch_distributed_examples_b200 import _native
    _native.set_instance(FakeNative())
    import pytorch_distributed_examples_b200 as pkg
    pkg.hvd.init()
    return pkg


def _mlp(seed):
    torch.manual_seed(seed)
    return nn.Sequential(nn.Linear(12, 7), nn.ReLU(), nn.Linear(7, 3))


def _data(step, rank, n=5):
    g = torch.Generator().manual_seed(1000 * step + rank)
    return torch.randn(n, 12, generator=g), torch.randint(0, 3, (n,), generator=g)


def _worker_optimizer(rank, world, port, q):
    pkg = _setup(rank, world, port)
    hvd = pkg.hvd
    assert hvd.size() == world and hvd.rank() == rank and hvd.local_rank() == rank
    model = _mlp(seed=10 + rank)                       # ranks start DIFFERENT ...
    opt = hvd.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9),
                                   named_parameters=model.named_parameters())
    assert isinstance(opt, torch.optim.SGD)
    hvd.broadcast_parameters(model.state_dict(), root_rank=0)     # ... and end equal to rank 0
    ref = _mlp(seed=10)
    for a, b in zip(model.parameters(), ref.parameters()):
        assert torch.equal(a, b)
    ropt = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9, foreach=False)
    for step in range(4):
        x, y = _data(step, rank)
        opt.zero_grad()
        F.cross_entropy(model(x), y).backward()
        opt.step()
        # oracle: average of every rank's gradient, then the plain optimizer step
        ropt.zero_grad()
        grads = None
        for r in range(world):
            xr, yr = _data(step, r)
            ref.zero_grad()
            F.cross_entropy(ref(xr), yr).backward()
            gs = [p.grad.clone() for p in ref.parameters()]
            grads = gs if grads is None else [a + b for a, b in zip(grads, gs)]
        for p, g in zip(ref.parameters(), grads):
            p.grad = g / world
        ropt.step()
    for a, b in zip(model.parameters(), ref.parameters()):
        torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)
    # explicit synchronize(): p.grad holds the averaged gradient, step() then must not average again
    x, y = _data(9, rank)
    opt.zero_grad(); F.cross_entropy(model(x), y).backward()
    local = [p.grad.clone() for p in model.parameters()]
    opt.synchronize()
    summed = [g.clone() for g in local]
    for g in summed:
        dist.all_reduce(g)
    for p, g in zip(model.parameters(), summed):
        torch.testing.assert_close(p.grad, g / world, rtol=1e-6, atol=1e-7)
    opt.step()
    # replicas bit-identical; sampler picks up hvd.size()/rank() (ref: mnist_horovod.py:41-42)
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    other = flat.clone(); dist.broadcast(other, src=0)
    assert torch.equal(flat, other)
    s = pkg.DistributedSampler(range(101), device="cpu")
    from torch.utils.data.distributed import DistributedSampler as TorchDS
    assert list(s) == list(TorchDS(range(101), num_replicas=world, rank=rank))
    assert hvd.broadcast_object({"a": rank}, root_rank=1) == {"a": 1}
    sd = opt.state_dict()
    assert "momentum_buffer" in sd["state"][0]
    # evaluate(): every rank ends with the GLOBAL (correct, total) of the sharded test set (one allreduce)
    shards = [[_data(50 + r * 3 + k, r, n=7 + k) for k in range(3)] for r in range(world)]
    with torch.no_grad():
        want_c = sum(int((model(x_).argmax(1) == y_).sum()) for sh in shards for x_, y_ in sh)
    want_t = sum(x_.shape[0] for sh in shards for x_, _ in sh)
    model.train()
    c_, t_, acc = pkg.evaluate(model, shards[rank], communicator=hvd._require())
    assert (c_, t_) == (want_c, want_t) and abs(acc - want_c / want_t) < 1e-12 and model.training
    q.put((rank, "ok"))
    dist.destroy_process_group()


def _worker_elastic(rank, world, port, q):
    pkg = _setup(rank, world, port)
    hvd = pkg.hvd
    lr = 0.01
    model = _mlp(seed=3 + rank)
    opt = hvd.DistributedOptimizer(torch.optim.AdamW(model.parameters(), lr=lr / math.sqrt(hvd.size())),
                                   named_parameters=model.named_parameters())
    log = []

    def on_state_reset():                                  # ref: horovod_mnist_elastic.py:80-82
        for g in opt.param_groups:
            g["lr"] = lr / math.sqrt(hvd.size())
        log.append(("reset", hvd.size()))

    @hvd.elastic.run
    def train(state):                                      # shape of ref: horovod_mnist_elastic.py:55-77
        batch_offset = state.batch
        for state.epoch in range(state.epoch, 2):
            for batch_idx in range(12):
                if batch_idx >= batch_offset:
                    x, y = _data(100 * state.epoch + batch_idx, hvd.rank())
                    opt.zero_grad(); F.cross_entropy(model(x), y).backward(); opt.step()
                    if state.epoch == 0 and batch_idx == 4 and hvd.size() == 2 and hvd.rank() == 0 and not state.shrunk:
                        hvd.elastic.request_resize([0])                     # 2 -> 1
                    if state.epoch == 1 and batch_idx == 2 and hvd.size() == 1:
                        hvd.elastic.request_resize([0, 1])                  # 1 -> 2
                    if state.batch % 3 == 0:
                        state.shrunk = state.shrunk or hvd.size() == 1
                        state.commit()
                        log.append(("commit", state.epoch, batch_idx, hvd.size()))
                    state.batch = batch_idx
            state.batch = 0
            batch_offset = 0

    state = hvd.elastic.TorchState(model, opt, batch=0, epoch=0, shrunk=False)
    state.register_reset_callbacks([on_state_reset])
    # commit / restore are bit-exact roll-backs
    state.commit()
    before = [p.detach().clone() for p in model.parameters()]
    with torch.no_grad():
        for p in model.parameters():
            p.add_(1.0)
    state.batch = 7
    state.restore()
    assert state.batch == 0 and all(torch.equal(a, b) for a, b in zip(before, model.parameters()))
    train(state)
    sizes = [e[1] for e in log if e[0] == "reset"]
    assert sizes == [1, 2] if rank == 0 else sizes == [2], (rank, log)      # rank 1 sat out the size-1 phase
    assert hvd.size() == 2
    assert abs(opt.param_groups[0]["lr"] - lr / math.sqrt(2)) < 1e-12
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    other = flat.clone(); dist.broadcast(other, src=0)
    assert torch.equal(flat, other), "after regrow + sync the replicas must be bit-identical"
    steps = opt._fused.groups[0].steps
    st = torch.tensor([steps]); dist.broadcast(st, src=0)
    assert int(st) == steps, "optimizer step counter must be synced too"
    q.put((rank, "ok"))
    dist.destroy_process_group()


def _run(worker, world=2):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
    alive = [p for p in procs if p.is_alive()]
    for p in alive:
        p.kill()
    assert not alive, "worker hung"
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    got = sorted(q.get(timeout=5) for _ in range(world))
    assert got == [(r, "ok") for r in range(world)]


def test_distributed_optimizer_and_broadcast_world2_gloo():
    _run(_worker_optimizer)


def test_elastic_commit_restore_resize_world2_gloo():
    _run(_worker_elastic)
