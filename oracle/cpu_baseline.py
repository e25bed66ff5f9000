"""The reference training loop timed on the host cores (reported baseline, never the optimisation target).

What runs is the reference's own configuration of the path:
  * `init_process_group("gloo")` + `OMP_NUM_THREADS`           ref: pytorch_elastic/mnist_ddp_elastic.py:22-27
  * `DDP(model, device_ids=None)` (stock CPU DDP: bucket allreduce over gloo)        ref: :58
  * loop body `Trainer._run_batch`: zero_grad / model(source) / criterion / backward / step   ref: :70-75
with either the Horovod scripts' CNN + SGD(+momentum) + F.nll_loss (ref: horovod/mnist_horovod.py:9-25,50,59-64 --
Horovod itself is not installable here, so its "average every p.grad, then step" is what stock DDP's bucket
allreduce does) or the DDP script's real model, `Model(hidden_layers=5, features=1024)` + Adam + CrossEntropyLoss
(ref: pytorch_elastic/mnist_ddp_elastic.py:133-159,172-174).  Models are the oracle restatements (net_oracle.OracleNet,
mlp_oracle.OracleMLP: bit-identical to the reference classes, checked in gen_golden.py); synthetic tensors are
pre-generated in RAM (MNIST cannot be downloaded).  `time_reference_ddp` launches W worker processes of this
module (one per rank, 127.0.0.1 rendezvous), times `steps` steps between barriers with time.perf_counter and
returns GLOBAL samples/s (W * per-rank batch / step time, max over ranks).
TEST/BENCH INFRASTRUCTURE ONLY: bench.py's ``cpu_baseline`` leg and ``--impl reference`` arm call this.
"""
from __future__ import annotations

import json
import os
import socket
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def cpu_model() -> str:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.lower().startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def _free_port() -> int:
    """A rendezvous port that is free NOW and lies below the kernel's ephemeral range (32768+): a port handed out by
    bind(0) sits inside that range and may be taken by an outgoing connection of the GPU job (NCCL / TCPStore sockets of
    8 ranks) before the CPU job's rank 0 listens on it -- seen once as EADDRINUSE at N = 8."""
    import random
    rng = random.Random(os.getpid() * 7919 + time.time_ns())
    for _ in range(200):
        port = rng.randrange(15000, 29000)
        with socket.socket() as s:
            try:
                s.bind(("127.0.0.1", port))
            except OSError:
                continue
            return port
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _build(model: str, opt_kw: dict):
    import torch
    import torch.nn.functional as F
    from . import mlp_oracle, net_oracle
    if model == "cnn":
        net = net_oracle.OracleNet()
        opt_kind = opt_kw.pop("kind", "sgd")
        crit = F.nll_loss
        shape = (1, 28, 28)
    elif model == "mlp":
        net = mlp_oracle.OracleMLP(5, 1024)
        opt_kind = opt_kw.pop("kind", "adam")
        crit = torch.nn.CrossEntropyLoss()
        shape = (1, 28, 28)
    else:
        raise ValueError(model)
    return net, opt_kind, crit, shape


def _worker(args):
    """One rank of the reference job (ref: mnist_ddp_elastic.py:22-27,58,70-75)."""
    import torch
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as DDP
    from . import net_oracle
    torch.set_num_threads(args["threads"])
    dist.init_process_group("gloo")                                   # env://, 127.0.0.1
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(0)                                              # same initial replica on every rank
    net, opt_kind, crit, shape = _build(args["model"], dict(args["opt"]))
    net.train()
    ddp = DDP(net, device_ids=None)
    opt_kw = {k: v for k, v in args["opt"].items() if k != "kind"}
    opt = net_oracle.make_optimizer(opt_kind, ddp.parameters(), **opt_kw)
    g = torch.Generator().manual_seed(1 + rank)                       # every rank its own shard
    x = torch.randn(args["batch"], *shape, generator=g)
    y = torch.randint(0, 10, (args["batch"],), generator=g)

    def run_batch():
        opt.zero_grad()
        loss = crit(ddp(x), y)
        loss.backward()
        opt.step()

    for _ in range(args["warmup"]):
        run_batch()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args["steps"]):
        run_batch()
    dist.barrier()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print("CPU_BASELINE_RESULT " + json.dumps(dict(
            seconds=float(t.item()), steps=args["steps"], batch=args["batch"], world=world,
            threads=torch.get_num_threads(), samples_per_s=world * args["batch"] * args["steps"] / float(t.item()),
            ms_per_step=float(t.item()) / args["steps"] * 1e3)), flush=True)
    dist.destroy_process_group()


def time_reference_ddp(world: int, batch: int, steps: int, warmup: int, threads_per_rank: int, model: str = "cnn",
                       opt: dict | None = None, timeout: float = 900.0) -> dict:
    """Launch the W-rank gloo job and return rank 0's result dict (global samples/s, ms/step, threads per rank).
    A rendezvous that loses its port to another process is retried on a fresh port."""
    last = None
    for _attempt in range(4):
        try:
            return _time_reference_ddp_once(world, batch, steps, warmup, threads_per_rank, model, opt, timeout)
        except RuntimeError as exc:
            last = exc
            if "EADDRINUSE" not in str(exc) and "address already in use" not in str(exc).lower():
                raise
    raise last


def _time_reference_ddp_once(world, batch, steps, warmup, threads_per_rank, model, opt, timeout) -> dict:
    opt = dict(opt or {})
    port = _free_port()
    spec = json.dumps(dict(model=model, batch=int(batch), steps=int(steps), warmup=int(warmup),
                           threads=int(threads_per_rank), opt=opt))
    procs = []
    for r in range(world):
        env = dict(os.environ)
        for k in list(env):                       # the GPU job's torchrun environment must not leak into the CPU job
            if k.startswith(("TORCHELASTIC", "GROUP_", "ROLE_")) or k in ("LOCAL_RANK", "LOCAL_WORLD_SIZE"):
                env.pop(k)
        env.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(r), WORLD_SIZE=str(world),
                   OMP_NUM_THREADS=str(threads_per_rank), MKL_NUM_THREADS=str(threads_per_rank), CUDA_VISIBLE_DEVICES="",
                   PYTHONPATH=ROOT + os.pathsep + env.get("PYTHONPATH", ""))
        procs.append(subprocess.Popen([sys.executable, "-m", "oracle.cpu_baseline", "--worker", spec], cwd=ROOT, env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    out0 = err0 = ""
    deadline = time.time() + timeout
    try:
        for r, p in enumerate(procs):
            o, e = p.communicate(timeout=max(1.0, deadline - time.time()))
            if r == 0:
                out0, err0 = o, e
            if p.returncode != 0:
                raise RuntimeError(f"cpu baseline rank {r} failed ({p.returncode}): {e[-2000:]}")
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    for line in out0.splitlines():
        if line.startswith("CPU_BASELINE_RESULT "):
            return json.loads(line[len("CPU_BASELINE_RESULT "):])
    raise RuntimeError("cpu baseline produced no result: " + err0[-2000:])


def bounded_reference_run(world: int, batch: int, steps: int, warmup: int, threads_per_rank: int, budget_s: float,
                          model: str = "cnn", opt: dict | None = None) -> dict:
    """`time_reference_ddp` on a BOUNDED sample of the workload: short probes size the per-rank slice of the
    batch so that `warmup + steps` steps fit in `budget_s` seconds (as large as fits: small slices under-state what the
    CPU reaches on the full batch); steps and warm-up are kept as asked.
    The result carries `sample_batch` (per-rank samples per step actually run) beside the full `batch`."""
    q = 64 if model == "cnn" else 16
    probe_b = min(batch, 128 if model == "cnn" else 32)
    sample_b = batch
    for _ in range(3):                      # the CPU's samples/s grows with the batch: re-probe at the size just chosen
        probe = time_reference_ddp(world, probe_b, 2, 1, threads_per_rank, model, opt)
        per_rank_rate = probe["samples_per_s"] / world
        need = batch * (steps + warmup) / max(per_rank_rate, 1e-9)
        if need <= budget_s:
            sample_b = batch
            break
        sample_b = min(batch, max(q, int(batch * budget_s / need) // q * q))
        if sample_b <= probe_b + probe_b // 2:
            break
        probe_b = sample_b
    r = time_reference_ddp(world, sample_b, steps, warmup, threads_per_rank, model, opt)
    r["sample_batch"] = sample_b
    r["full_batch"] = batch
    r["ms_per_full_step"] = r["ms_per_step"] * batch / sample_b
    return r


if __name__ == "__main__":
    if len(sys.argv) >= 3 and sys.argv[1] == "--worker":
        _worker(json.loads(sys.argv[2]))
    else:       # by hand: python -m oracle.cpu_baseline W B steps threads [cnn|mlp]
        W, B, K, T = (int(v) for v in sys.argv[1:5])
        print(json.dumps(time_reference_ddp(W, B, K, 2, T, sys.argv[5] if len(sys.argv) > 5 else "cnn",
                                            dict(kind="sgd", lr=0.01, momentum=0.9) if len(sys.argv) <= 5 or sys.argv[5] == "cnn"
                                            else dict(kind="adam", lr=1e-3))))
