"""Step time of the DDP script's MLP (Model(hidden_layers=5, features=1024) + CrossEntropyLoss + Adam,
ref: pytorch_elastic/mnist_ddp_elastic.py:172-174): native tensor-core path vs torch/cuBLAS on the same GPU.
    python tools/mlp_bench.py [--batch 4096] [--steps 30]
Prints one JSON line.  17,364,992 MAC/sample => 34.7 MFLOP/sample fwd+bwd."""
import argparse
import json
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
MAC_PER_SAMPLE = 17_364_992


def timed(fn, steps, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return statistics.median(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=30)
    args = ap.parse_args()
    import pytorch_distributed_examples_b200 as pkg
    from oracle import mlp_oracle
    B = args.batch
    ds = pkg.SyntheticMNIST(length=1 << 20, seed=1)
    x, y = ds.batch(base_index=0, count=B)
    x = x.reshape(B, -1).contiguous()
    out = {"batch": B, "model": "Model(hidden_layers=5, features=1024), 6,062,090 params", "flop_per_step": 2.0 * MAC_PER_SAMPLE * B}
    torch.manual_seed(0)
    sd = mlp_oracle.OracleMLP(5, 1024).state_dict()

    # native: fused loss+grads, then the fused Adam update
    model = pkg.Model(5, 1024); model.load_state_dict(sd); model = model.cuda()
    opt = pkg.FusedOptimizer(torch.optim.Adam(model.parameters(), lr=1e-3))
    def native_fb():
        model.loss_and_grads(x, y)
    def native_step():
        model.loss_and_grads(x, y)
        opt.step()
    ms = timed(native_fb, args.steps)
    out["native_fwd_bwd_ms"] = ms
    out["native_fwd_bwd_tflops_fp32_equiv"] = out["flop_per_step"] / (ms * 1e-3) / 1e12
    out["native_step_ms"] = timed(native_step, args.steps)
    # module path (autograd + torch CrossEntropyLoss), the drop-in loop body
    crit = torch.nn.CrossEntropyLoss()
    def native_module_step():
        opt.zero_grad()
        crit(model(x), y).backward()
        opt.step()
    out["native_module_step_ms"] = timed(native_module_step, args.steps)

    # torch on the same GPU: strict fp32 (the reference's arithmetic) and with TF32 tensor cores allowed (NOT fp32-accurate)
    for name, allow in (("torch_fp32", False), ("torch_tf32_allowed", True)):
        torch.backends.cuda.matmul.allow_tf32 = allow
        ref = mlp_oracle.OracleMLP(5, 1024); ref.load_state_dict(sd); ref = ref.cuda()
        ropt = torch.optim.Adam(ref.parameters(), lr=1e-3, fused=True)
        def ref_fb():
            ref.zero_grad(set_to_none=True)
            crit(ref(x), y).backward()
        def ref_step():
            ropt.zero_grad()
            crit(ref(x), y).backward()
            ropt.step()
        ms = timed(ref_fb, args.steps)
        out[name + "_fwd_bwd_ms"] = ms
        out[name + "_fwd_bwd_tflops"] = out["flop_per_step"] / (ms * 1e-3) / 1e12
        out[name + "_step_ms"] = timed(ref_step, args.steps)
    torch.backends.cuda.matmul.allow_tf32 = False
    out["speedup_vs_torch_fp32_fwd_bwd"] = out["torch_fp32_fwd_bwd_ms"] / out["native_fwd_bwd_ms"]
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
