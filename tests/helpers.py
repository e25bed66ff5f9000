"""Shared helpers for the parity tests."""
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

# BASELINE.json north_star: "per-step loss and parameters within 1e-5 relative in fp32"
RTOL = 1e-5
# gradients/activations are sums of O(100..1e6) products: compare against the tensor's scale
def assert_close_scaled(actual, expected, rtol=RTOL, name=""):
    actual = actual.detach().float().cpu()
    expected = expected.detach().float().cpu()
    assert actual.shape == expected.shape, (name, actual.shape, expected.shape)
    scale = max(expected.abs().max().item(), 1e-30)
    err = (actual - expected).abs().max().item()
    assert err <= rtol * scale, f"{name}: max|err|={err:.3e} > {rtol:g} * scale {scale:.3e} (rel {err / scale:.3e})"
    return err / scale


def load_golden(name):
    return torch.load(os.path.join(GOLDEN, name), map_location="cpu", weights_only=False)


def arena_from_state_dict(nat, sd, device="cuda"):
    off, sz = nat.param_layout()
    names = ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")
    arena = torch.zeros(21848, dtype=torch.float32, device=device)
    for n, o, s in zip(names, off, sz):
        arena[o:o + s] = sd[n].reshape(-1).to(device)
    return arena


def split_arena(nat, arena):
    off, sz = nat.param_layout()
    names = ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")
    shapes = ((10, 1, 5, 5), (10,), (20, 10, 5, 5), (20,), (50, 320), (50,), (10, 50), (10,))
    return {n: arena[o:o + s].view(sh) for n, o, s, sh in zip(names, off, sz, shapes)}
