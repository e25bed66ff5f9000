"""The DDP script's own configuration in situ (ref: pytorch_elastic/mnist_ddp_elastic.py:58,70-75,172-174):
Model(hidden_layers=5, features=1024) + Adam + CrossEntropyLoss under stock DistributedDataParallel, one rank per GPU.
    torchrun --nproc-per-node=N tools/mlp_ddp_bench.py [--batch 4096]
Variants: torch modules with strict-fp32 cuBLAS (the reference's arithmetic on this GPU) vs pkg.Model (tensor-core
3xTF32 kernels) with our bucket allreduce hook and the fused Adam update.  One JSON line, step time = max over ranks."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
from torch.nn.parallel import DistributedDataParallel as DDP

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=8)
    args = ap.parse_args()
    import pytorch_distributed_examples_b200 as pkg
    hvd = pkg.hvd
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    hvd.init()
    W, R = hvd.size(), hvd.rank()
    torch.backends.cuda.matmul.allow_tf32 = False
    ds = pkg.SyntheticMNIST(length=1 << 22, seed=3)
    x, y = ds.batch(base_index=R * args.batch, count=args.batch)
    crit = torch.nn.CrossEntropyLoss()
    out = {"n_gpus": W, "per_gpu_batch": args.batch, "model": "Model(hidden_layers=5, features=1024) + Adam(lr=1e-3)"}

    def torch_model():
        layers = [torch.nn.Flatten(), torch.nn.Linear(784, 1024), torch.nn.ReLU()]
        for _ in range(5):
            layers += [torch.nn.Linear(1024, 1024), torch.nn.ReLU()]
        return torch.nn.Sequential(*layers, torch.nn.Linear(1024, 10))

    for variant in ("torch_fp32_stock_ddp", "native_allreduce_hook", "native_allreduce_hook_fused_adam"):
      try:
          torch.manual_seed(0)
          model = (torch_model() if variant.startswith("torch") else pkg.Model(5, 1024)).cuda()
          ddp = DDP(model, device_ids=[local], gradient_as_bucket_view=True)
          if variant.startswith("native"):
              ddp.register_comm_hook(pkg.HookState(), pkg.allreduce_hook)
          opt = torch.optim.Adam(ddp.parameters(), lr=1e-3, fused=True)
          if variant.endswith("fused_adam"):
              opt = pkg.FusedOptimizer(torch.optim.Adam(ddp.parameters(), lr=1e-3))

          def step():
              opt.zero_grad()
              loss = crit(ddp(x), y)
              loss.backward()
              opt.step()
              return loss

          first_loss = float(step().detach())          # identical across variants: same seed, same batch
          for _ in range(args.warmup - 1):
              step()
          dist.barrier(); torch.cuda.synchronize()
          e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
          e0.record()
          for _ in range(args.steps):
              loss = step()
          e1.record(); dist.barrier(); torch.cuda.synchronize()
          t = torch.tensor([e0.elapsed_time(e1) / args.steps], device="cuda", dtype=torch.float64)
          dist.all_reduce(t, op=dist.ReduceOp.MAX)
          out[variant] = {"ms_per_step": t.item(), "samples_per_s": W * args.batch / (t.item() * 1e-3), "first_step_loss": first_loss,
                          "last_loss": float(loss.detach()),
                          "note": "the same batch is replayed every step, so late losses sit near zero and diverge chaotically"}
          del ddp, opt, model
          torch.cuda.empty_cache()
      except Exception as exc:                      # keep the other variants' numbers
        out[variant] = {"error": repr(exc)}
        dist.barrier()
    best = min(v["ms_per_step"] for k, v in out.items() if k.startswith("native") and isinstance(v, dict) and "ms_per_step" in v)
    out["speedup_native_vs_torch_fp32"] = out["torch_fp32_stock_ddp"]["ms_per_step"] / best
    if R == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
