"""In-situ check of the DDP comm hooks on a foreign model at scale (BASELINE configs[4]):
ResNet-50 (the architecture of ref: rpc/model_parallel_ResNet50.py:94-101,126-132 — its two nn.Sequential shards are
torchvision's ResNet-50 cut in half), stock torch DDP, bf16 autocast, synthetic 224x224, 25.6 M parameters.
conv/BN kernels stay with cuDNN (out of scope); what is ours is the bucket path: allreduce through libddp_step.so
(allreduce_hook) or allreduce + fused SGD-momentum update inside the hook (FusedUpdateHook).
    torchrun --nproc-per-node=N tools/resnet_ddp_bench.py [--batch 32] [--steps 20]
Prints one JSON line: step time per variant (max over ranks, CUDA events) and the parameter agreement between them."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    args = ap.parse_args()
    import torchvision
    import pytorch_distributed_examples_b200 as pkg
    from torch.nn.parallel import DistributedDataParallel as DDP
    hvd = pkg.hvd
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    hvd.init()
    W, R = hvd.size(), hvd.rank()
    g = torch.Generator(device="cuda").manual_seed(100 + R)
    x = torch.randn(args.batch, 3, 224, 224, device="cuda", generator=g)
    y = torch.randint(0, 1000, (args.batch,), device="cuda", generator=g)
    crit = torch.nn.CrossEntropyLoss()
    out = {"n_gpus": W, "per_gpu_batch": args.batch, "model": "resnet50 (25,557,032 params)", "autocast": "bf16"}
    finals = {}
    sym_kw = dict(symmetric_bytes=40 << 20)      # DDP's fp32 buckets are <= ~30 MiB (25 MiB cap + one tensor)
    for variant in ("stock_ddp", "allreduce_hook", "fused_update_hook", "allreduce_hook_sym", "fused_update_hook_sym"):
        torch.manual_seed(0)
        model = torchvision.models.resnet50(num_classes=1000).cuda()
        ddp = DDP(model, device_ids=[local], gradient_as_bucket_view=True)
        opt = torch.optim.SGD(ddp.parameters(), lr=0.01, momentum=0.9)
        hook_state = None
        if variant == "allreduce_hook":
            ddp.register_comm_hook(pkg.HookState(), pkg.allreduce_hook)
        elif variant == "fused_update_hook":
            ddp.register_comm_hook(pkg.FusedUpdateHook.State(opt), pkg.FusedUpdateHook.hook)
        elif variant == "allreduce_hook_sym":
            hook_state = pkg.HookState(**sym_kw)
            ddp.register_comm_hook(hook_state, pkg.allreduce_hook)
        elif variant == "fused_update_hook_sym":
            hook_state = pkg.FusedUpdateHook.State(opt, **sym_kw)
            ddp.register_comm_hook(hook_state, pkg.FusedUpdateHook.hook)

        def step():
            opt.zero_grad(set_to_none=False)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = crit(ddp(x), y)
            loss.backward()
            if not variant.startswith("fused_update_hook"):
                opt.step()
            return loss

        for _ in range(args.warmup):
            step()
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            loss = step()
        e1.record(); dist.barrier(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / args.steps], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        finals[variant] = torch.cat([p.detach().float().reshape(-1) for p in model.parameters()])
        out[variant] = {"ms_per_step": t.item(), "images_per_s": W * args.batch / (t.item() * 1e-3), "loss": float(loss.detach())}
        if hook_state is not None:
            out[variant]["symmetric_bucket_calls"] = hook_state.sym_calls
            out[variant]["hook_calls"] = hook_state.calls
        del ddp, opt, model
        torch.cuda.empty_cache()
    ref = finals["stock_ddp"]
    for v in ("allreduce_hook", "fused_update_hook", "allreduce_hook_sym", "fused_update_hook_sym"):
        out[v]["max_param_diff_vs_stock"] = float((finals[v] - ref).abs().max())
    out["param_scale"] = float(ref.abs().max())
    if R == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
