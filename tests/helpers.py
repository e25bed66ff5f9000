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


def saved_routing(workspace, batch):
    """Our forward's saved activations out of the caller-owned workspace (layout: csrc/common.cuh carve_workspace):
    a1 [B,1440] f32, idx1 u8, a2 [B,320] f32, idx2 u8, h [B,50] f32 -- the ReLU masks and pool argmaxes the
    backward kernels route gradients by."""
    a256 = lambda v: (v + 255) & ~255
    B = batch
    o = 0
    a1 = workspace[o:o + B * 1440 * 4].view(torch.float32).view(B, 1440); o += a256(B * 1440 * 4)
    idx1 = workspace[o:o + B * 1440].view(B, 1440); o += a256(B * 1440)
    a2 = workspace[o:o + B * 320 * 4].view(torch.float32).view(B, 320); o += a256(B * 320 * 4)
    idx2 = workspace[o:o + B * 320].view(B, 320); o += a256(B * 320)
    h = workspace[o:o + B * 50 * 4].view(torch.float32).view(B, 50)
    return dict(a1=a1.cpu(), idx1=(idx1 & 3).cpu().long(), a2=a2.cpu(), idx2=(idx2 & 3).cpu().long(), h=h.cpu())


def count_routing_flips(ours, ref, mask2d=None, mask1d=None, rel=1e-5):
    """Disagreements between OUR gradient routes and the fp64 oracle's (oracle/net_oracle.routing).
    Returns (flips, bad): `flips` = routes that differ where a gradient actually flows; `bad` = those of them whose
    fp64 margin (pool: best minus runner-up; ReLU: |pre-activation|) exceeds rel * the layer's scale, i.e. NOT a
    near-tie.  A correct fp32 kernel has bad == 0 at any batch size; a kernel that mis-routes for a systematic
    reason does not."""
    flips = bad = 0
    for L, pre_k, idx_k, mar_k, act_k, sc_k in ((1, "pre1", "idx1", "margin1", "a1", "scale1"),
                                               (2, "pre2", "idx2", "margin2", "a2", "scale2")):
        pre, scale = ref[pre_k], ref[sc_k]
        live_ref, live_ours = pre > 0, ours[act_k] > 0
        relu_flip = live_ref != live_ours
        pool_flip = (ref[idx_k] != ours[idx_k]) & live_ref & live_ours      # a dead cell routes nothing either way
        if L == 2 and mask2d is not None:                                    # dropped channels route nothing
            keep = (mask2d.view(-1, 20, 1) > 0).expand(-1, 20, 16).reshape(pre.shape)
            relu_flip &= keep; pool_flip &= keep
        flips += int(relu_flip.sum()) + int(pool_flip.sum())
        bad += int((relu_flip & (pre.abs() > rel * scale)).sum()) + int((pool_flip & (ref[mar_k] > rel * scale)).sum())
    z, scale = ref["pre_fc1"], ref["scale_fc1"]
    h_live = ours["h"] > 0
    ref_live = z > 0
    if mask1d is not None:
        ref_live = ref_live & (mask1d > 0)
    f = h_live != ref_live
    flips += int(f.sum()); bad += int((f & (z.abs() > rel * scale)).sum())
    return flips, bad
