"""The reference training loop timed on the host cores (reported baseline, never the optimisation target).

Loop body = ref: horovod/mnist_horovod.py:59-64 (zero_grad / model(data) / F.nll_loss / backward / step) with
the reference's own CNN (restated in net_oracle.OracleNet, bit-identical to the reference class — checked in
gen_golden.py) under the installed torch CPU kernels, synthetic tensors pre-generated in RAM.
TEST/BENCH INFRASTRUCTURE ONLY: bench.py's ``cpu_baseline`` leg and ``--impl reference`` arm call this.
"""
from __future__ import annotations

import time

import torch
import torch.nn.functional as F

from . import net_oracle


def time_reference_loop(batch: int, steps: int, warmup: int, opt_kind: str = "sgd", threads: int | None = None,
                        seed: int = 0, **opt_kw):
    """Returns dict(samples_per_s, ms_per_step, threads, steps, batch)."""
    if threads:
        torch.set_num_threads(threads)
    torch.manual_seed(seed)
    net = net_oracle.OracleNet()
    net.train()
    opt = net_oracle.make_optimizer(opt_kind, net.parameters(), **opt_kw)
    g = torch.Generator().manual_seed(seed + 1)
    x = torch.randn(batch, 1, 28, 28, generator=g)
    y = torch.randint(0, 10, (batch,), generator=g)

    def one():
        opt.zero_grad()
        out = net(x)
        loss = F.nll_loss(out, y)
        loss.backward()
        opt.step()
        return loss

    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt = time.perf_counter() - t0
    return dict(samples_per_s=batch * steps / dt, ms_per_step=dt / steps * 1e3, threads=torch.get_num_threads(),
                steps=steps, batch=batch)
