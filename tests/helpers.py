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


def assert_grad_close(actual, g32, g64, batch, rtol=RTOL, flips=4, name=""):
    """Parity of a batch-MEAN gradient of a ReLU/max-pool network.
    Such gradients are discontinuous in the activations: when an activation sits within rounding noise of a ReLU
    threshold or a pooling tie, two correct implementations route that ONE sample's gradient differently and the
    mean moves by about scale/B — e.g. 3.5e-4 of the scale for conv1.weight at B=4096, which is also how far
    torch's own fp32 CPU run is from its fp64 run on that input.  Tolerance, per element:
        rtol*scale (arithmetic)  +  |g32 - g64| (the oracle's own routing ambiguity on this input)
                                 +  flips * scale / B (a few routing flips of our own).
    For B <= 64 the last term is dropped by the callers (flips=0): the golden tests hold the plain 1e-5 bar."""
    a = actual.detach().double().cpu()
    g32 = g32.detach().double().cpu()
    g64 = g64.detach().double().cpu()
    scale = max(g64.abs().max().item(), 1e-30)
    tol = rtol * scale + (g32 - g64).abs() + flips * scale / batch
    err = torch.minimum((a - g32).abs(), (a - g64).abs())
    bad = err > tol
    assert not bool(bad.any()), (f"{name}: {int(bad.sum())} elements off, worst {err.max().item():.3e} "
                                 f"(scale {scale:.3e}, allowance {rtol * scale + flips * scale / batch:.3e})")
    return err.max().item() / scale
