#!/usr/bin/env python
"""bench.py — MNIST-CNN data-parallel training step on N B200s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl reference]
N>1 is launched by torchrun (one rank per GPU).  Rank 0 prints ONE JSON line.

A "step" is one pass of the hot path over one synthetic batch: forward + NLL, backward, bucket allreduce
(N>1), fused average + SGD-momentum update — the reference loop of mnist_horovod.py:59-64 on its CNN.
`value` is whole-job samples/s with inputs resident in HBM (rotating through more input bytes than the L2 holds),
`e2e` is the same step driven from pinned HOST buffers (H2D copies + loss read-back inside the timed region).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_SAMPLE = 2_595_000          # fwd 961,000 + bwd 1,634,000
KERNEL_MACS = {                      # algorithmic MACs per sample of each kernel (dense formulation)
    "conv1_fwd": 144_000, "conv2_fwd": 320_000, "head_fwd": 16_500, "head_bwd": 33_000,
    "conv2_dgrad": 320_000, "conv2_wgrad": 320_000, "conv1_wgrad": 144_000, "reduce_partials": 0,
}
METRIC = "mnist_cnn_train_samples_per_sec"
_JSON_OUT = sys.stdout
OPT = dict(kind="sgd", lr=0.01, momentum=0.9)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(hbm=p["hbm_gbs"], tflops=p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                    tflops_burst=p["bf16_tflops"], source="measured")
    return dict(hbm=6650.0, tflops=1400.0, tflops_burst=1590.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.idx = [], None, gpu_index
        self.t_lo = self.t_hi = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
        except Exception:
            self.proc = None
            return
        def pump():
            for line in self.proc.stdout:
                self.rows.append((time.time(), line.strip()))
        threading.Thread(target=pump, daemon=True).start()

    def window(self, lo, hi):
        self.t_lo, self.t_hi = lo, hi

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if self.t_lo is None or self.t_lo - 0.05 <= t <= self.t_hi + 0.15] or \
               [r for _, r in self.rows]
        sm, mx, reasons = [], [], set()
        for r in rows:
            f = [c.strip() for c in r.split(",")]
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except Exception:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference(n_ranks, B, steps, warmup, budget_s):
    """The reference's W-rank CPU/gloo loop on this box's host cores (oracle/cpu_baseline.py): stock CPU DDP around the
    reference CNN, SGD(lr=0.01, momentum=0.9), per-rank batch B, W = n_ranks processes.  Two thread settings:
    all host threads (cores // W per rank; the arm's `value`) and the reference-faithful OMP_NUM_THREADS=1
    (ref: pytorch_elastic/mnist_ddp_elastic.py:27; torchrun's own default).  Each on a bounded sample."""
    from oracle import cpu_baseline
    cores = cpu_baseline.host_cores()
    per_rank = max(1, cores // n_ranks)
    opt = dict(kind=OPT["kind"], lr=OPT["lr"], momentum=OPT["momentum"])
    best = cpu_baseline.bounded_reference_run(n_ranks, B, steps, warmup, per_rank, budget_s, "cnn", opt)
    faithful = best if per_rank == 1 else cpu_baseline.bounded_reference_run(n_ranks, B, steps, warmup, 1, budget_s, "cnn", opt)
    def describe(r):
        return (f"{r['steps']} timed steps of a {r['sample_batch']}-sample-per-rank slice of the {B}-sample-per-rank step, "
                f"{n_ranks}-rank gloo DDP job (stock CPU DDP, reference loop body), {r['threads']} thread(s) per rank")
    return dict(value=best["samples_per_s"], unit="samples/s", cores=per_rank * n_ranks, kind="port", sample=describe(best),
                ms_per_full_step=best["ms_per_full_step"], ranks=n_ranks, threads_per_rank=per_rank, host_cores=cores,
                cpu_model=cpu_baseline.cpu_model(),
                reference_faithful_1_thread_per_rank={"value": faithful["samples_per_s"], "unit": "samples/s",
                                                      "cores": n_ranks, "sample": describe(faithful),
                                                      "ms_per_full_step": faithful["ms_per_full_step"]})


def reference_arm(args, world, rank):
    """--impl reference: the reference's CPU implementation of the step on the host cores.  Rank 0 alone launches and
    prints it; what it launches is the reference's own configuration -- an N-rank gloo DDP job (N = --gpus) of the
    reference loop, per-rank batch as in the GPU arm -- so the two arms' `config` are the same workload."""
    if rank != 0:
        return
    B, N = args.batch, args.gpus
    r = cpu_reference(N, B, args.steps, args.warmup, budget_s=70.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": "samples/s", "n_gpus": N,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_full_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": workload_config(B, N),
        "cpu_baseline": r,
        "e2e": {"value": r["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _JSON_OUT.write(json.dumps(line) + "\n"); _JSON_OUT.flush()


def workload_config(B, n):
    return {"workload": f"MNIST CNN Net (mnist_horovod.py) fused fwd/bwd/optimizer step, synthetic 28x28, "
                        f"batch {B}/GPU, fp32, SGD lr=0.01 momentum=0.9" + (", per-step bucket allreduce" if n > 1 else ""),
            "per_gpu_batch": B, "global_batch": B * n, "optimizer": "sgd(lr=0.01,momentum=0.9)",
            "parallelism": f"dp{n}", "l2": "inputs rotate through a resident pool larger than the 126 MB L2"}


def step_latency_stats(pkg, nat, B, momentum, steps=1000, warmup=50, slots=4):
    """Per-step latency distribution of the graph-replayed step at batch B on one GPU (BASELINE configs[1]/[2] sizes):
    one CUDA event per step boundary, consecutive differences -> median / p99 / mean."""
    import torch
    torch.manual_seed(0)
    model = pkg.Net().cuda().train()
    opt = pkg.FusedOptimizer(torch.optim.SGD(model.parameters(), lr=OPT["lr"], momentum=momentum))
    st = pkg.TrainStep(model, opt, B, slots=slots)
    ds = pkg.SyntheticMNIST(length=1 << 30, seed=1234)
    for s_ in range(slots):
        nat.synth_batch(ds.seed, None, s_ * B, B, st.xs[s_], st.ys[s_])
    for i in range(warmup):
        st.run_slot(i % slots)
    torch.cuda.synchronize()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    evs[0].record()
    for i in range(steps):
        st.run_slot(i % slots)
        evs[i + 1].record()
    torch.cuda.synchronize()
    d = sorted(evs[i].elapsed_time(evs[i + 1]) * 1e3 for i in range(steps))
    return {"batch": B, "optimizer": f"sgd(lr=0.01,momentum={momentum})", "steps": steps,
            "us_median": d[len(d) // 2], "us_p99": d[min(len(d) - 1, int(len(d) * 0.99))], "us_mean": sum(d) / len(d),
            "samples_per_s_at_median": B / (d[len(d) // 2] * 1e-6)}


def module_path_stats(pkg, nat, B, steps=200, warmup=20):
    """The five-line loop the reference scripts contain, through the drop-in seam (ref: horovod/mnist_horovod.py:59-64):
    optimizer.zero_grad(); output = model(data); loss = F.nll_loss(output, target); loss.backward(); optimizer.step()
    with `Net` (autograd.Function) + `hvd.DistributedOptimizer` -- eager launches, no CUDA graph."""
    import torch
    import torch.nn.functional as F
    torch.manual_seed(0)
    model = pkg.Net().cuda().train()
    opt = pkg.hvd.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=OPT["lr"], momentum=OPT["momentum"]),
                                       named_parameters=model.named_parameters())
    ds = pkg.SyntheticMNIST(length=1 << 30, seed=1234)
    batches = [ds.batch(base_index=i * B, count=B) for i in range(4)]
    def one(i):
        data, target = batches[i % 4]
        opt.zero_grad()
        output = model(data)
        loss = F.nll_loss(output, target)
        loss.backward()
        opt.step()
        return loss
    for i in range(warmup):
        one(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        loss = one(i)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / steps
    return {"batch": B, "us_per_step": us, "samples_per_s": B / (us * 1e-6), "steps": steps, "final_loss": float(loss.item())}


def bucket_allreduce_stats(comm, N, reps=20):
    """ResNet-50-sized gradient bucket (BASELINE configs[4]: 25,557,032 parameters) over NVLink at N ranks: ncclAllReduce
    and the in-switch two-shot allreduce (ddp_sym_allreduce), fp32 (102 MB) and bf16 (51 MB).  Collective: every rank calls it.
    busbw = (S / t) * 2 (N - 1) / N (nccl-tests convention); t = max over ranks of the mean over `reps` back-to-back calls."""
    import torch
    import torch.distributed as dist
    out = {}
    n = 25_557_032 // 8 * 8
    have_sym = comm.enable_symmetric(n * 4 + 4096)
    for dtype, name in ((torch.float32, "fp32"), (torch.bfloat16, "bf16")):
        S = n * (4 if dtype == torch.float32 else 2)
        buf = torch.randn(n, device="cuda").to(dtype)
        def timed(fn):
            for _ in range(3):
                fn()
            torch.cuda.synchronize(); dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                fn()
            b.record(); torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b) / reps], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item()) * 1e-3
        t = timed(lambda: comm.allreduce_sum_(buf))
        out[f"nccl_{name}"] = {"bytes": S, "us": t * 1e6, "busbw_GB/s": S / t * 2 * (N - 1) / N / 1e9,
                               "frac_of_900": S / t * 2 * (N - 1) / N / 900e9}
        if have_sym:
            comm.sym_view(dtype, n).copy_(buf)
            t = timed(lambda: comm.sym_allreduce_(dtype, n, 0, 1.0 / N))
            out[f"sym_two_shot_{name}"] = {"bytes": S, "us": t * 1e6, "busbw_GB/s": S / t * 2 * (N - 1) / N / 1e9,
                                           "frac_of_900": S / t * 2 * (N - 1) / N / 900e9,
                                           "mode": "multimem (in-switch)" if comm.sym["multicast"] else "peer load/store"}
        del buf
    if not have_sym:
        out["sym_two_shot"] = {"unavailable": comm.sym_error}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--batch", type=int, default=4096, help="per-GPU batch (BASELINE configs[1]: 4096)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="skip cpu_baseline / HBM-kernel extras")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries ONE JSON line and nothing else: libraries that write to file descriptor 1 (NCCL's version banner,
    # torchrun children) are sent to stderr for the duration of the run; the line itself goes to the saved descriptor
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        reference_arm(args, world, rank)
        return
    args.warmup = max(args.warmup, 3)

    # keep stdout to the ONE JSON line: NCCL_DEBUG=VERSION (this image's default) makes NCCL print its banner there
    if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"
    import torch
    import torch.distributed as dist
    import pytorch_distributed_examples_b200 as pkg
    from pytorch_distributed_examples_b200 import _native
    hvd = pkg.hvd

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    hvd.init()
    nat = _native.get()
    comm = hvd._require()
    N, B, K, W = world, args.batch, args.steps, args.warmup
    peaks = measured_peaks()

    torch.manual_seed(0)
    model = pkg.Net().cuda().train()
    opt = hvd.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=OPT["lr"], momentum=OPT["momentum"]),
                                   named_parameters=model.named_parameters())
    hvd.broadcast_parameters(model.state_dict(), root_rank=0)
    # resident input pool larger than L2 (126 MB): each slot is one batch of this rank's shard
    slot_bytes = B * (784 * 4 + 8)
    slots = max(2, min(64, -(-160_000_000 // slot_bytes)))
    step = pkg.TrainStep(model, opt, B, communicator=comm if N > 1 else None, slots=slots)
    ds = pkg.SyntheticMNIST(length=1 << 30, seed=1234)
    for s in range(slots):
        nat.synth_batch(ds.seed, None, (rank * slots + s) * B, B, step.xs[s], step.ys[s])
    torch.cuda.synchronize()

    def barrier():
        if N > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if N == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- device-resident timed region ----------------
    for i in range(W):
        step.run_slot(i % slots)
    clocks = ClockSampler(local)
    clocks.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_lo = time.time()
    e0.record()
    for i in range(K):
        step.run_slot(i % slots)
    e1.record()
    barrier()
    t_hi = time.time()
    clocks.window(t_lo, t_hi)
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    clock_info = clocks.stop()
    loss_final = float(step.loss.item())
    value = N * B * K / (ms_total * 1e-3)

    # ---------------- end-to-end: pinned host buffers -> H2D -> step -> loss D2H ----------------
    # Every step's batch is copied from pinned host memory inside the timed region and its loss is read back on
    # the host; the copy of batch i+1 is prefetched on a copy stream while batch i computes (the pin_memory +
    # non_blocking pattern of the reference's DataLoader, ref: mnist_ddp_elastic.py:183-187), and the loss of
    # step i is read one step late (post_loss/read_loss) so the host never stalls the launch queue.
    # The host batches are raw uint8 pixels (MNIST's storage format); ToTensor+Normalize runs on the device.
    host_slots = min(slots, 8)
    hx, hy = [], []
    for s in range(host_slots):
        px, yy = ds.batch_u8(base_index=(rank * slots + s) * B, count=B)
        hx.append(px.cpu().pin_memory()); hy.append(yy.cpu().pin_memory())
    Ke = max(10, min(K, 400))
    torch.cuda.synchronize()
    step.stage_u8(0, hx[0], hy[0])
    for i in range(3):
        step.stage_u8((i + 1) % 2, hx[(i + 1) % host_slots], hy[(i + 1) % host_slots])
        step.run_slot(i % 2).item()
    barrier()
    e0.record()
    prev = None
    for i in range(3, Ke + 3):
        step.stage_u8((i + 1) % 2, hx[(i + 1) % host_slots], hy[(i + 1) % host_slots])
        step.run_slot(i % 2)
        step.post_loss(i % 2)
        if prev is not None:
            loss_host = step.read_loss(prev)                                  # D2H read of step i-1's loss
        prev = i % 2
    loss_host = step.read_loss(prev)
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    e2e_value = N * B * Ke / (ms_e2e * 1e-3)
    assert loss_host == loss_host, "e2e loss is NaN"

    line = {
        "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": N, "steps": K, "warmup": W,
        "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32", "data": "synthetic", "config": workload_config(B, N), "clocks": clock_info,
        "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": B * 784 + B * 8,
                "input_format": "uint8 pixels [B,784] + int64 labels from pinned host memory; ToTensor+Normalize on device",
                "d2h_bytes_per_step": 4, "steps": Ke, "ms_per_step": ms_e2e / Ke},
        "gpu_launches": step.kernel_launches_per_step * K,
        "final_loss": loss_final,
    }
    if N > 1:
        # data parallelism's invariant, checked on the final state of the timed job: every replica holds bit-identical
        # parameters (min and max over ranks of an exact integer checksum of the parameter bits agree)
        bits = model._arena.view(torch.int32).to(torch.int64)
        chk = torch.stack([bits.sum(), (bits * torch.arange(1, bits.numel() + 1, device=bits.device)).sum()])
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        line["replicas_bit_identical"] = bool(torch.equal(lo, hi))
        assert line["replicas_bit_identical"], "replicas diverged: the gradient exchange is broken"
    if N > 1:    # which exchange ran: the one-shot NVLink peer-memory allreduce fused with the update, or ncclAllReduce + fused update
        line["collective"] = "nvlink-peer-memory one-shot allreduce fused with the update" if step.use_peer \
            else "ncclAllReduce + fused update"

    # ---------------- roofline of the dominant kernel (rank 0) ----------------
    if rank == 0:
        nat.prof_enable(True)
        acc = {k: 0.0 for k in _native.PROF_KERNELS}
        reps = 30
        eager = pkg.TrainStep(model, opt, B, communicator=None, use_graph=False, slots=1)
        eager.xs[0].copy_(step.xs[0]); eager.ys[0].copy_(step.ys[0])
        upd_ms = 0.0
        for i in range(reps + 3):
            nat.net_forward(model._arena, step.xs[i % slots], step.ys[i % slots], None, None, 1, i, True, step.ws,
                            step.logp, out_loss=step.loss, scratch=step.scratch)
            nat.net_backward(model._arena, step.xs[i % slots], step.logp, None, step.ys[i % slots], 1.0, True, step.ws,
                             step.scratch, model._grad_arena)
            if i >= 3:
                for k, v in nat.prof_read().items():
                    acc[k] += v
        nat.prof_enable(False)
        per_kernel_us = {k: v / reps * 1e3 for k, v in acc.items()}
        top = max(per_kernel_us, key=per_kernel_us.get)
        dur_s = per_kernel_us[top] * 1e-6
        flops = 2.0 * KERNEL_MACS[top] * B
        achieved = flops / dur_s / 1e12
        sm_mhz = clock_info.get("sm_mhz") or 1900.0
        ffma_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12
        traffic, traffic_src = None, None
        for tname in ("r02_ncu_traffic_b4096.json", "r01_ncu_traffic_b4096.json"):
            tpath = os.path.join(ROOT, "profiles", tname)
            if B == 4096 and os.path.exists(tpath):      # DRAM bytes per launch: NOT measured in this run
                traffic = json.load(open(tpath)).get(top)
                traffic_src = f"from committed ncu --set full capture profiles/{tname}, not measured in this run"
                if traffic is not None:
                    break
        # the conv2 backward kernels visit only the live (channel, pool cell) entries: executed MACs = 250 per entry
        a256 = lambda v: (v + 255) & ~255
        o_a2 = a256(B * 1440 * 4) + a256(B * 1440)
        live_per_sample = float((step.ws[o_a2:o_a2 + B * 320 * 4].view(torch.float32) > 0).float().sum().item()) / B
        executed_macs = dict(KERNEL_MACS)
        if os.environ.get("DDP_BWD_DENSE", "0") != "1":
            executed_macs["conv2_dgrad"] = executed_macs["conv2_wgrad"] = 250.0 * live_per_sample
        line["roofline"] = {"kernel": top, "bound": "tensor", "achieved": achieved, "peak": peaks["tflops"],
                            "unit": "TFLOP/s", "frac": achieved / peaks["tflops"], "traffic": traffic,
                            "traffic_source": traffic_src,
                            "peak_source": f"MEASURED_PEAKS.json bf16 sustained ({peaks['source']})",
                            "note": "achieved = dense-equivalent algorithmic FLOPs (SURVEY 8(d) M3) / measured duration; "
                                    "fp32 kernels (1e-5 parity bar rules out single-pass TF32/BF16): fraction of the fp32 "
                                    "FFMA peak beside it; executed_* count the multiply-adds the kernel really issues "
                                    "(the sparse conv2 backward skips structural zeros of the pool / Dropout2d / ReLU gradient)",
                            "fp32_ffma_peak_tflops": ffma_peak, "frac_of_fp32_ffma_peak": achieved / ffma_peak,
                            "executed_tflops": 2.0 * executed_macs[top] * B / dur_s / 1e12,
                            "live_conv2_entries_per_sample": live_per_sample,
                            "kernel_us": per_kernel_us,
                            "kernel_dense_equiv_tflops": {k: 2.0 * KERNEL_MACS[k] * B / max(v, 1e-9) / 1e6 for k, v in per_kernel_us.items()},
                            "kernel_executed_tflops": {k: 2.0 * executed_macs[k] * B / max(v, 1e-9) / 1e6 for k, v in per_kernel_us.items()},
                            "step_share": per_kernel_us[top] / max(sum(per_kernel_us.values()), 1e-9)}
        step_flops = FLOP_PER_SAMPLE * B
        line["step_tflops"] = step_flops / (ms_total / K * 1e-3) / 1e12

    # ---------------- collective extra (all ranks): ResNet-50-sized bucket allreduce over NVLink ----------------
    bucket_stats = None
    if N > 1 and not args.no_extras:
        try:
            bucket_stats = bucket_allreduce_stats(comm, N)
        except Exception as exc:
            bucket_stats = {"error": repr(exc)}

    # ---------------- extras: HBM-bound kernels, other BASELINE configs, CPU baselines (rank 0) ----------------
    if rank == 0 and not args.no_extras:
        extra = {}
        if bucket_stats is not None:
            extra[f"resnet50_bucket_allreduce_n{N}"] = bucket_stats
        n = 25_557_032 // 4 * 4     # ResNet-50-sized bucket (BASELINE configs[4])
        g = torch.randn(n, device="cuda"); p = torch.randn(n, device="cuda"); s1 = torch.zeros(n, device="cuda")
        h = _native.Hyper(); h.lr = 0.01; h.momentum = 0.9
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        def timed(fn, reps=10):
            ts = []
            for _ in range(reps + 2):
                flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); fn(); b.record(); torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
            return statistics.median(ts[2:]) * 1e-3
        t = timed(lambda: nat.fused_update(_native.OPT_SGD, g, p, s1, None, n, 0.125, h, step=2))
        extra["fused_update_sgd_momentum_25.6M"] = {"GB/s": 20.0 * n / t / 1e9, "frac_hbm": 20.0 * n / t / 1e9 / peaks["hbm"],
                                                     "us": t * 1e6, "bytes": 20 * n}
        s2 = torch.zeros(n, device="cuda"); h2 = _native.Hyper(); h2.lr = 1e-3; h2.beta1 = 0.9; h2.beta2 = 0.999; h2.eps = 1e-8; h2.weight_decay = 0.01
        t = timed(lambda: nat.fused_update(_native.OPT_ADAMW, g, p, s1, s2, n, 0.125, h2, step=2))
        extra["fused_update_adamw_25.6M"] = {"GB/s": 28.0 * n / t / 1e9, "frac_hbm": 28.0 * n / t / 1e9 / peaks["hbm"],
                                              "us": t * 1e6, "bytes": 28 * n}
        from pytorch_distributed_examples_b200.bucket import BucketTable, FlatLayout, flatten_into
        chunks = list(torch.randn(n, device="cuda").split(n // 161 + 1))
        layout = FlatLayout.for_tensors(chunks); bucket = torch.empty(layout.total, device="cuda"); table = BucketTable(torch.device("cuda"))
        flatten_into(bucket, chunks, layout, table)
        t = timed(lambda: flatten_into(bucket, chunks, layout, table))
        extra["flatten_bucket_161_tensors_102MB"] = {"GB/s": 8.0 * n / t / 1e9, "frac_hbm": 8.0 * n / t / 1e9 / peaks["hbm"], "us": t * 1e6}
        xb = torch.empty(65536, 1, 28, 28, device="cuda"); yb = torch.empty(65536, dtype=torch.int64, device="cuda")
        t = timed(lambda: nat.synth_batch(1, None, 0, 65536, xb, yb))
        extra["synth_batch_65536"] = {"GB/s": 65536 * 3144 / t / 1e9, "frac_hbm": 65536 * 3144 / t / 1e9 / peaks["hbm"], "us": t * 1e6}
        # the DDP script's own model: native tensor-core MLP step vs torch/cuBLAS strict fp32, same GPU
        try:
            Bm = 4096
            xm, ym = ds.batch(base_index=0, count=Bm)
            xm = xm.reshape(Bm, -1).contiguous()
            torch.manual_seed(0)
            mlp = pkg.Model(5, 1024).cuda()
            t_nat = timed(lambda: mlp.loss_and_grads(xm, ym), reps=10)
            torch.backends.cuda.matmul.allow_tf32 = False      # the reference's arithmetic: strict fp32 matmuls
            layers = [torch.nn.Linear(784, 1024), torch.nn.ReLU()]
            for _ in range(5):
                layers += [torch.nn.Linear(1024, 1024), torch.nn.ReLU()]
            tm = torch.nn.Sequential(*layers, torch.nn.Linear(1024, 10)).cuda()
            with torch.no_grad():
                for dst, src in zip(tm.parameters(), mlp.parameters()):
                    dst.copy_(src)
            def torch_fb():
                tm.zero_grad(set_to_none=True)
                torch.nn.functional.cross_entropy(tm(xm), ym).backward()
            t_ref = timed(torch_fb, reps=10)
            flop = 2.0 * 17_364_992 * Bm
            extra["mlp_5x1024_fwd_bwd_b4096"] = {"native_us": t_nat * 1e6, "native_tflops_fp32_equiv": flop / t_nat / 1e12,
                                                 "torch_fp32_cublas_us": t_ref * 1e6, "speedup": t_ref / t_nat,
                                                 "frac_of_tf32_tensor_peak": 3 * flop / t_nat / 1e12 / (peaks["tflops"] / 2),
                                                 "note": "3xTF32 tcgen05 GEMMs (3 tensor-core products per fp32 product); "
                                                         "tf32 peak taken as half the measured bf16 peak; timings include an L2 flush"}
            del mlp, tm, xm, ym
        except Exception as exc:       # the MLP row is an extra: never lose the headline line over it
            extra["mlp_5x1024_fwd_bwd_b4096"] = {"error": repr(exc)}
        if N == 1:
            try:      # BASELINE configs[1] / configs[2] sizes and the reference-exact optimizer (SGD, momentum 0: mnist_horovod.py:50)
                extra["step_latency_b64_sgd_momentum"] = step_latency_stats(pkg, nat, 64, 0.9)
                extra["step_latency_b512_sgd_momentum"] = step_latency_stats(pkg, nat, 512, 0.9)
                extra["step_latency_b64_sgd_mu0"] = step_latency_stats(pkg, nat, 64, 0.0)
                extra["step_latency_b4096_sgd_mu0"] = step_latency_stats(pkg, nat, 4096, 0.0, steps=300)
                extra["step_latency_b4096_sgd_momentum"] = step_latency_stats(pkg, nat, 4096, 0.9, steps=300)
                extra["module_path_b4096"] = module_path_stats(pkg, nat, 4096)
                extra["module_path_b64"] = module_path_stats(pkg, nat, 64)
            except Exception as exc:
                extra["latency_extras_error"] = repr(exc)
            try:      # the DDP script's real configuration on the host cores: Model(5, 1024) + Adam + CrossEntropyLoss (:172-174)
                from oracle import cpu_baseline
                rm = cpu_baseline.bounded_reference_run(1, 4096, 2, 1, cpu_baseline.host_cores(), 10.0, "mlp", dict(kind="adam", lr=1e-3))
                extra["mlp_5x1024_adam_cpu_reference"] = {"samples_per_s": rm["samples_per_s"], "ms_per_4096_step": rm["ms_per_full_step"],
                                                          "threads": rm["threads"], "sample": f"{rm['steps']} steps of a {rm['sample_batch']}-sample slice, 1-rank gloo DDP"}
            except Exception as exc:
                extra["mlp_5x1024_adam_cpu_reference"] = {"error": repr(exc)}
        line["extra"] = extra
        del g, p, s1, s2, flush, chunks, bucket, xb, yb
        pass
    if rank == 0 and not args.no_extras:
        # the reference's W-rank CPU/gloo loop on this box's host cores, W = N, bounded sample
        try:
            line["cpu_baseline"] = cpu_reference(N, B, steps=4, warmup=1, budget_s=12.0)
        except Exception as exc:          # a host-side hiccup of the CPU job must not cost the measured GPU line
            line["cpu_baseline"] = {"value": None, "unit": "samples/s", "cores": None, "kind": "port",
                                    "sample": "failed", "error": repr(exc)[:400]}
    if rank == 0:
        _JSON_OUT.write(json.dumps(line) + "\n"); _JSON_OUT.flush()
    if N > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
