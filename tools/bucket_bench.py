"""Bucket path at ResNet-50 scale (BASELINE configs[4]): flatten -> ncclAllReduce(sum) over
NVLink -> fused (1/N)*g + SGD-momentum update, 25,557,032 parameters (fp32 102 MB / bf16 51 MB bucket).
Run under torchrun, one rank per GPU:  torchrun --nproc-per-node=N tools/bucket_bench.py
Prints one JSON line with the allreduce bus bandwidth (nccl-tests convention busbw = S/t * 2(N-1)/N) and
the per-stage times, timed with CUDA events, max over ranks."""
import json
import os
import statistics
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import pytorch_distributed_examples_b200 as pkg
    from pytorch_distributed_examples_b200 import _native
    from pytorch_distributed_examples_b200.bucket import BucketTable, FlatLayout, flatten_into
    hvd = pkg.hvd
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    hvd.init()
    nat = _native.get()
    c = hvd._require()
    N, R = c.world, c.rank
    P = 25_557_032
    # ResNet-50-like tensor size mix: 161 tensors
    g = torch.Generator().manual_seed(0)
    cuts = sorted(torch.randint(1, P, (160,), generator=g).tolist())
    sizes = [b - a for a, b in zip([0] + cuts, cuts + [P])]
    out = {"n_gpus": N, "params": P, "tensors": len(sizes)}

    def timed(fn, reps=20, warm=5):
        ts = []
        for i in range(reps + warm):
            if N > 1:
                dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b)], device="cuda", dtype=torch.float64)
            if N > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if i >= warm:
                ts.append(t.item())
        return statistics.median(ts) * 1e-3

    for dtype, name in ((torch.float32, "fp32"), (torch.bfloat16, "bf16")):
        grads = [torch.randn(n, device="cuda").to(dtype) for n in sizes]
        layout = FlatLayout.for_tensors(grads)
        bucket = torch.zeros(layout.total, dtype=dtype, device="cuda")
        table = BucketTable(torch.device("cuda"))
        params = torch.randn(layout.total, device="cuda"); mom = torch.zeros_like(params)
        h = _native.Hyper(); h.lr = 0.1; h.momentum = 0.9
        S = bucket.numel() * bucket.element_size()
        t_flat = timed(lambda: flatten_into(bucket, grads, layout, table))
        t_ar = timed(lambda: c.allreduce_sum_(bucket))
        t_upd = timed(lambda: nat.fused_update(_native.OPT_SGD, bucket, params, mom, None, layout.total, 1.0 / N, h, step=2))

        def whole():
            flatten_into(bucket, grads, layout, table)
            c.allreduce_sum_(bucket)
            nat.fused_update(_native.OPT_SGD, bucket, params, mom, None, layout.total, 1.0 / N, h, step=2)
        t_all = timed(whole)
        busbw = S / t_ar * 2 * (N - 1) / N / 1e9 if N > 1 else None
        sym = {}
        if N > 1 and c.enable_symmetric(S):
            n_ar = -(-layout.total // 8) * 8          # whole 16-byte vectors; the pad elements are zero
            c.sym_view(dtype, n_ar).zero_()
            v = c.sym_view(dtype, layout.total)
            v.copy_(bucket)
            saved = c.sym["multicast"]
            avail = c.sym["multicast_available"]
            def back_to_back(fn, iters=20, reps=5):
                """nccl-tests convention: `iters` launches back to back, one event pair around them, per-launch mean."""
                ts = []
                for _ in range(reps + 1):
                    dist.barrier(); torch.cuda.synchronize()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    for _ in range(iters):
                        fn()
                    b.record(); torch.cuda.synchronize()
                    t = torch.tensor([a.elapsed_time(b) / iters], device="cuda", dtype=torch.float64)
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    ts.append(t.item())
                return statistics.median(ts[1:]) * 1e-3

            def row(t):
                return {"us": t * 1e6, "busbw_GBps": S / t * 2 * (N - 1) / N / 1e9, "busbw_frac_of_900": S / t * 2 * (N - 1) / N / 1e9 / 900}
            bucket.zero_()                           # sums stay finite over the back-to-back launches
            t = back_to_back(lambda: c.allreduce_sum_(bucket))
            sym["nccl_back_to_back"] = row(t)
            for mode, mc in ((("multicast", avail),) if avail else ()) + (("peer", 0),):
                c.sym["multicast"] = mc
                variants = ((0, 0), (64, 1), (48, 2), (96, 2), (64, 4)) if mc else ((0, 0),)
                for ctas, un in variants:
                    tag = f"{mode}_ctas{ctas or 'default'}_unroll{un or 'default'}"
                    fn = lambda: c.sym_allreduce_(dtype, n_ar, 0, 1.0 / N, max_ctas=ctas, unroll=un)
                    sym[tag] = {"single_shot": row(timed(fn, reps=10, warm=3)), "back_to_back": row(back_to_back(fn))}
            c.sym["multicast"] = saved

            def whole_sym():
                flatten_into(v, grads, layout, table)
                c.sym_allreduce_(dtype, n_ar, 0, 1.0)
                nat.fused_update(_native.OPT_SGD, v, params, mom, None, layout.total, 1.0 / N, h, step=2)
            sym["flatten+allreduce+update_us"] = timed(whole_sym) * 1e6
        elif N > 1:
            sym = {"unavailable": c.sym_error}
        out[name] = {"bucket_bytes": S, "flatten_us": t_flat * 1e6, "flatten_GBps": 2 * S / t_flat / 1e9,
                     "allreduce_us": t_ar * 1e6, "allreduce_algbw_GBps": S / t_ar / 1e9, "allreduce_busbw_GBps": busbw,
                     "busbw_frac_of_900": busbw / 900 if busbw else None,
                     "fused_update_us": t_upd * 1e6,
                     "fused_update_GBps": (S + 16 * layout.total) / t_upd / 1e9,
                     "flatten+allreduce+update_us": t_all * 1e6, "symmetric_two_shot": sym}
        del grads, bucket, params, mom
    # the CNN's 87 KB bucket: latency
    small = torch.zeros(21848, device="cuda")
    out["cnn_bucket_87KB_allreduce_us"] = timed(lambda: c.allreduce_sum_(small), reps=50) * 1e6
    if R == 0:
        print(json.dumps(out), flush=True)
    if N > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
