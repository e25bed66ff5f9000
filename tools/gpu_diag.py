"""Kernel-by-kernel diagnostics on a GPU box (never stops at the first failure).  Writes gpurun_out/diag.txt."""
import os
import sys
import time
import traceback

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
OUT = open(os.path.join(ROOT, "gpurun_out", "diag.txt"), "w")


def say(*a):
    s = " ".join(str(x) for x in a)
    print(s, flush=True)
    OUT.write(s + "\n"); OUT.flush()


def rel(a, b):
    a = a.detach().float().cpu(); b = b.detach().float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def section(name, fn):
    say(f"=== {name}")
    try:
        fn()
    except Exception:
        say("EXCEPTION", traceback.format_exc())


def main():
    from helpers import arena_from_state_dict, load_golden, split_arena
    from oracle import net_oracle, sampler_oracle, synth_oracle
    from pytorch_distributed_examples_b200 import _native
    nat = _native.get()
    nat.bind_device(0)
    say(torch.cuda.get_device_name(0), torch.__version__)

    def fwdbwd(name):
        fx = load_golden(name)
        B = fx["x"].shape[0]
        arena = arena_from_state_dict(nat, fx["state_dict"])
        ws = torch.zeros(nat.workspace_bytes(B), dtype=torch.uint8, device="cuda")
        sc = torch.zeros(nat.scratch_bytes(B), dtype=torch.uint8, device="cuda")
        logp = torch.zeros(B, 10, device="cuda"); loss = torch.zeros(1, device="cuda")
        x, y = fx["x"].cuda(), fx["y"].cuda()
        nat.net_forward(arena, x, y, fx["mask2d"].cuda(), fx["mask1d"].cuda(), 0, 0, True, ws, logp, out_loss=loss, scratch=sc)
        torch.cuda.synchronize()
        # intermediate activations vs oracle
        o = net_oracle.OracleNet(); o.load_state_dict(fx["state_dict"])
        import torch.nn.functional as F
        with torch.no_grad():
            a1 = F.relu(F.max_pool2d(o.conv1(fx["x"]), 2))
            c2 = o.conv2(a1) * (fx["mask2d"].view(-1, 20, 1, 1) * 2)
            a2 = F.relu(F.max_pool2d(c2, 2)).view(-1, 320)
            h = F.relu(o.fc1(a2)) * (fx["mask1d"] * 2)
        al = lambda n: (n + 255) // 256 * 256
        off = 0
        a1g = ws[off:off + B * 1440 * 4].view(torch.float32).view(B, 10, 12, 12); off += al(B * 1440 * 4) + al(B * 1440)
        a2g = ws[off:off + B * 320 * 4].view(torch.float32).view(B, 320); off += al(B * 320 * 4) + al(B * 320)
        hg = ws[off:off + B * 50 * 4].view(torch.float32).view(B, 50)
        say(name, "a1", rel(a1g, a1), "a2", rel(a2g, a2), "h", rel(hg, h), "logp", rel(logp, fx["logp"]),
            "loss", loss.item(), fx["loss"].item())
        bucket = torch.full((21848,), float("nan"), device="cuda")
        nat.net_backward(arena, x, logp, None, y, 1.0, True, ws, sc, bucket)
        torch.cuda.synchronize()
        grads = split_arena(nat, bucket)
        say(name, "grads", {k: f"{rel(grads[k], g):.2e}" for k, g in fx["grads"].items()})

    for nm in ("fwdbwd_b1.pt", "fwdbwd_b8.pt", "fwdbwd_b48.pt", "fwdbwd_b64.pt"):
        section(nm, lambda nm=nm: fwdbwd(nm))

    def sampler():
        for n, w, r in [(10, 2, 1), (1000, 3, 2), (60000, 8, 5)]:
            ns = nat.sampler_num_samples(n, w, False)
            perm = torch.empty(n, dtype=torch.int32, device="cuda"); out = torch.empty(ns, dtype=torch.int64, device="cuda")
            t = time.time(); nat.sampler_indices(n, w, r, True, 0, 1, False, perm, out); torch.cuda.synchronize(); dt = time.time() - t
            ref = sampler_oracle.shard_indices(n, w, r, True, 0, 1)
            say("sampler", n, w, r, "exact" if np.array_equal(out.cpu().numpy(), ref) else "MISMATCH", f"{dt * 1e3:.2f} ms")
    section("sampler", sampler)

    def synth():
        x = torch.empty(16, 1, 28, 28, device="cuda"); y = torch.empty(16, dtype=torch.int64, device="cuda")
        nat.synth_batch(1234, None, 5, 16, x, y)
        xo, yo = synth_oracle.synth_batch(1234, np.arange(5, 21))
        say("synth x exact", np.array_equal(x.cpu().numpy(), xo), "y exact", np.array_equal(y.cpu().numpy(), yo))
    section("synth", synth)

    def update():
        n = 21848
        p_ref = torch.nn.Parameter(torch.randn(n)); ref = torch.optim.AdamW([p_ref], lr=0.01, foreach=False)
        p = p_ref.detach().clone().cuda(); s1 = torch.zeros(n, device="cuda"); s2 = torch.zeros(n, device="cuda")
        h = _native.Hyper(); h.lr = 0.01; h.weight_decay = 0.01; h.beta1 = 0.9; h.beta2 = 0.999; h.eps = 1e-8
        for step in range(1, 4):
            g = torch.randn(n); p_ref.grad = g; ref.step()
            nat.fused_update(_native.OPT_ADAMW, g.cuda(), p, s1, s2, n, 1.0, h, step=step)
        say("adamw rel", rel(p, p_ref))
    section("update", update)

    def timing():
        B = 4096
        torch.manual_seed(0)
        sd = net_oracle.OracleNet().state_dict()
        arena = arena_from_state_dict(nat, sd)
        x = torch.randn(B, 1, 28, 28, device="cuda"); y = torch.randint(0, 10, (B,), device="cuda")
        ws = torch.zeros(nat.workspace_bytes(B), dtype=torch.uint8, device="cuda")
        sc = torch.zeros(nat.scratch_bytes(B), dtype=torch.uint8, device="cuda")
        logp = torch.zeros(B, 10, device="cuda"); loss = torch.zeros(1, device="cuda")
        bucket = torch.zeros(21848, device="cuda")
        for _ in range(3):
            nat.net_forward(arena, x, y, None, None, 1, 0, True, ws, logp, out_loss=loss, scratch=sc)
            nat.net_backward(arena, x, logp, None, y, 1.0, True, ws, sc, bucket)
        torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        for _ in range(20):
            nat.net_forward(arena, x, y, None, None, 1, 0, True, ws, logp, out_loss=loss, scratch=sc)
        e[1].record()
        for _ in range(20):
            nat.net_backward(arena, x, logp, None, y, 1.0, True, ws, sc, bucket)
        e[2].record(); torch.cuda.synchronize()
        say(f"B={B} fwd {e[0].elapsed_time(e[1]) / 20 * 1e3:.1f} us  bwd {e[1].elapsed_time(e[2]) / 20 * 1e3:.1f} us")
    section("timing", timing)


if __name__ == "__main__":
    main()
