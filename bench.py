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


def reference_arm(args, world, rank):
    """--impl reference: the reference's CPU implementation of the step on the host cores (oracle port)."""
    if rank != 0:
        return
    import torch
    from oracle import cpu_baseline
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    B = args.batch
    probe = cpu_baseline.time_reference_loop(min(B, 256), 2, 1, OPT["kind"], lr=OPT["lr"], momentum=OPT["momentum"])
    est_full = B / probe["samples_per_s"]                       # seconds per full-size step
    budget = 150.0
    sample_b = B
    if est_full * (args.steps + args.warmup) > budget:          # bound the per-step sample, keep K and W
        sample_b = max(64, int(B * budget / (est_full * (args.steps + args.warmup))) // 64 * 64)
    r = cpu_baseline.time_reference_loop(sample_b, args.steps, args.warmup, OPT["kind"], lr=OPT["lr"],
                                         momentum=OPT["momentum"])
    sample = (f"{args.steps} timed steps of a {sample_b}-sample slice of the {B}-sample step, "
              f"torch {torch.__version__} CPU, {r['threads']} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": r["samples_per_s"], "unit": "samples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"] * (B / sample_b),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": workload_config(B, args.gpus),
        "cpu_baseline": {"value": r["samples_per_s"], "unit": "samples/s", "cores": r["threads"], "kind": "port",
                         "sample": sample},
        "e2e": {"value": r["samples_per_s"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(B, n):
    return {"workload": f"MNIST CNN Net (mnist_horovod.py) fused fwd/bwd/optimizer step, synthetic 28x28, "
                        f"batch {B}/GPU, fp32, SGD lr=0.01 momentum=0.9" + (", per-step bucket allreduce" if n > 1 else ""),
            "per_gpu_batch": B, "global_batch": B * n, "optimizer": "sgd(lr=0.01,momentum=0.9)",
            "parallelism": f"dp{n}", "l2": "inputs rotate through a resident pool larger than the 126 MB L2"}


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
    if N > 1:    # which exchange ran: the one-shot NVLink peer-memory allreduce fused with the update, or ncclAllReduce + fused update
        line["config"]["collective"] = "nvlink-peer-memory one-shot allreduce fused with the update" if step.use_peer \
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
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "r01_ncu_traffic_b4096.json")
        if B == 4096 and os.path.exists(tpath):          # DRAM bytes per launch from the committed ncu --set full capture
            traffic = json.load(open(tpath)).get(top)
        line["roofline"] = {"kernel": top, "bound": "tensor", "achieved": achieved, "peak": peaks["tflops"],
                            "unit": "TFLOP/s", "frac": achieved / peaks["tflops"], "traffic": traffic,
                            "peak_source": f"MEASURED_PEAKS.json bf16 sustained ({peaks['source']})",
                            "note": "fp32 FFMA kernel (1e-5 parity bar rules out single-pass TF32/BF16); "
                                    "fraction of the fp32 FFMA peak given beside it",
                            "fp32_ffma_peak_tflops": ffma_peak, "frac_of_fp32_ffma_peak": achieved / ffma_peak,
                            "kernel_us": per_kernel_us,
                            "step_share": per_kernel_us[top] / max(sum(per_kernel_us.values()), 1e-9)}
        step_flops = FLOP_PER_SAMPLE * B
        line["step_tflops"] = step_flops / (ms_total / K * 1e-3) / 1e12

    # ---------------- extras: HBM-bound kernels + CPU baseline (rank 0, N=1) ----------------
    if rank == 0 and not args.no_extras:
        extra = {}
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
        line["extra"] = extra
        del g, p, s1, s2, flush, chunks, bucket, xb, yb
        if N == 1:
            from oracle import cpu_baseline
            cores = os.cpu_count() or 1
            probe = cpu_baseline.time_reference_loop(256, 2, 1, OPT["kind"], threads=cores, lr=OPT["lr"], momentum=OPT["momentum"])
            nsteps = max(2, min(40, int(15.0 * probe["samples_per_s"] / B)))
            r = cpu_baseline.time_reference_loop(B, nsteps, 1, OPT["kind"], threads=cores, lr=OPT["lr"], momentum=OPT["momentum"])
            line["cpu_baseline"] = {"value": r["samples_per_s"], "unit": "samples/s", "cores": r["threads"], "kind": "port",
                                    "sample": f"{nsteps} steps of the same {B}-sample step (reference loop, torch CPU, "
                                              f"{r['threads']} threads, {r['ms_per_step']:.1f} ms/step)"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if N > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
