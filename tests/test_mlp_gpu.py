"""MLP path: tf32 hi/lo split, the tcgen05 3xTF32 GEMM, and the full forward/backward of the DDP
script's Model against the oracle (reference Model under torch CPU fp32/fp64) — all through the C-ABI."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_split_tf32_planes(native):
    torch.manual_seed(0)
    src = torch.randn(100, 76, device="cuda") * torch.logspace(-3, 3, 76, device="cuda")
    hi, lo, hiT, loT = native.split_tf32(src, transposed=True)
    assert torch.all((hi.view(torch.int32) & 0x1FFF) == 0) and torch.all((lo.view(torch.int32) & 0x1FFF) == 0), "planes must be tf32-exact"
    err = ((hi.double() + lo.double()) - src.double()).abs()
    assert torch.all(err <= src.double().abs() * 2.0 ** -21), "hi + lo must reproduce the value to ~22 bits"
    assert torch.equal(hiT, hi.t().contiguous()) and torch.equal(loT, lo.t().contiguous())


@pytest.mark.parametrize("M,N,K", [(128, 128, 32), (256, 256, 64), (200, 136, 100), (200, 136, 1000), (1000, 784, 512),
                                   (2048, 1024, 784), (1024, 1024, 4096), (4, 12, 8), (1024, 784, 260)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
def test_gemm_3xtf32_matches_fp64(native, M, N, K, a_mn, b_mn):
    """Every operand-layout combination the step uses (forward K/K, data gradient K/MN, weight gradient MN/MN)."""
    if (a_mn and M % 4) or (b_mn and N % 4):
        pytest.skip("MN-major operands need a contiguous dimension that is a multiple of 4")
    torch.manual_seed(M + N + K)
    a = torch.randn(M, K, device="cuda")
    b = torch.randn(N, K, device="cuda")
    a_st = a.t().contiguous() if a_mn else a            # [K][M] when MN-major
    b_st = b.t().contiguous() if b_mn else b
    # MN-major tiles may read (never use) a few bytes past the last row: keep the planes inside a larger allocation
    def planes(t):
        hi, lo = native.split_tf32(t)
        pad_hi = torch.zeros(t.numel() + 64, device="cuda"); pad_lo = torch.zeros(t.numel() + 64, device="cuda")
        pad_hi[:t.numel()].copy_(hi.reshape(-1)); pad_lo[:t.numel()].copy_(lo.reshape(-1))
        return pad_hi[:t.numel()].view(t.shape), pad_lo[:t.numel()].view(t.shape)
    a_hi, a_lo = planes(a_st)
    b_hi, b_lo = planes(b_st)
    full = torch.full((M, N + 4), 7.0, device="cuda")                  # ldc > N: padding columns must stay untouched
    out = full[:, :N]
    native.gemm_3xtf32(a_hi, a_lo, b_hi, b_lo, out=out, a_mn_major=a_mn, b_mn_major=b_mn)
    ref64 = a.double() @ b.double().t()
    scale = (a.double().abs() @ b.double().abs().t())                  # per-element magnitude of the summed products
    err = (out.double() - ref64).abs()
    assert torch.all(err <= 1e-6 * scale + 1e-30), f"max err/scale {float((err / scale).max()):.3e}"
    err32 = ((a @ b.t()).double() - ref64).abs().max().item()         # cuBLAS fp32 on the same inputs, for scale
    assert err.max().item() <= 4 * err32 + 1e-6 * ref64.abs().max().item()
    assert torch.all(full[:, N:] == 7.0)


def _grad_close(name, got, g32, g64, rtol=1e-5):
    """|got - g64| <= rtol*scale + the fp32 oracle's own distance from fp64 (its rounding is not ours to beat)."""
    scale = g64.abs().max().item() + 1e-30
    err = (got.double().cpu() - g64).abs().max().item()
    own = (g32.double() - g64).abs().max().item()
    assert err <= rtol * scale + 2 * own, f"{name}: err {err:.3e}, scale {scale:.3e}, fp32-oracle err {own:.3e}"


@pytest.mark.parametrize("name", ["mlp_h1_f128.pt", "mlp_h2_f64.pt"])
def test_mlp_matches_reference_golden(native, name):
    """Forward, loss, gradients and 3 Adam steps of the reference Model/Trainer._run_batch (golden fixtures written
    by oracle/gen_golden.py from the reference module itself)."""
    import pytorch_distributed_examples_b200 as pkg
    from helpers import load_golden
    from oracle import mlp_oracle
    fx = load_golden(name)
    H, Fe = fx["hidden_layers"], fx["features"]
    model = pkg.Model(hidden_layers=H, features=Fe)
    model.load_state_dict(fx["state_dict0"])
    model = model.cuda()
    x0, y0 = fx["xs"][0].cuda(), fx["ys"][0].cuda()
    logits = model(x0)
    loss = torch.nn.CrossEntropyLoss()(logits, y0)
    loss.backward()
    _, _, g64 = mlp_oracle.forward_backward(fx["state_dict0"], fx["xs"][0], fx["ys"][0], H, Fe, dtype=torch.float64)
    assert (logits.detach().cpu() - fx["logits"]).abs().max().item() <= 1e-5 * fx["logits"].abs().max().item()
    assert abs(loss.item() - fx["loss"].item()) <= 1e-5 * abs(fx["loss"].item())
    for k, p in model.named_parameters():
        _grad_close(k, p.grad, fx["grads"][k], g64[k])
    # fused path: same loss, same gradients
    fused = pkg.Model(hidden_layers=H, features=Fe); fused.load_state_dict(fx["state_dict0"]); fused = fused.cuda()
    loss_f = fused.loss_and_grads(x0, y0)
    assert abs(loss_f.item() - fx["loss"].item()) <= 1e-5 * abs(fx["loss"].item())
    for (k, p), (_, q) in zip(model.named_parameters(), fused.named_parameters()):
        assert (p.grad - q.grad).abs().max().item() <= 2e-6 * p.grad.abs().max().item() + 1e-12, k
    # the reference loop body with Adam (ref :70-75, :173)
    model.zero_grad(set_to_none=True)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    crit = torch.nn.CrossEntropyLoss()
    losses = []
    for k in range(fx["xs"].shape[0]):
        opt.zero_grad()
        out = model(fx["xs"][k].cuda())
        l = crit(out, fx["ys"][k].cuda())
        l.backward()
        opt.step()
        losses.append(l.item())
    np.testing.assert_allclose(losses, fx["losses"].numpy(), rtol=2e-5)
    steps = fx["xs"].shape[0]
    for k, v in model.state_dict().items():
        ref = fx["state_dict"][k]
        d = (v.cpu() - ref).abs()
        tol = 1e-5 * ref.abs().max().item() + 1e-3 * 1e-3            # Adam: |g| ~ eps elements move by O(lr) on 1-ulp grad changes
        assert (d <= tol).float().mean().item() >= 0.98 and d.max().item() <= 0.25 * 1e-3 * steps, k


def test_mlp_full_size_against_oracle(native):
    """The script's real configuration, Model(hidden_layers=5, features=1024) (ref :172), batch 256.
    A ReLU unit whose pre-activation is within rounding of zero can fall on the other side than in the oracle
    ("flip"): that is a legitimate fp32 outcome (torch's own GPU fp32 flips 1-2 units per layer at B=4096,
    tools/mlp_diag.py) but moves gradients by O(1/B).  So the gradients are checked twice: (a) tightly, 1e-5,
    against an fp64 backward that uses OUR forward's ReLU masks (rebuilt here from the same GEMM building
    blocks); (b) against the plain oracle with an allowance of 4 flips, like the CNN tests."""
    import pytorch_distributed_examples_b200 as pkg
    from oracle import mlp_oracle, synth_oracle
    torch.manual_seed(3)
    ref = mlp_oracle.OracleMLP(5, 1024)
    sd = {k: v.detach().clone() for k, v in ref.state_dict().items()}
    B = 256
    xn, yn = synth_oracle.synth_batch(1234, np.arange(B))
    x, y = torch.from_numpy(xn), torch.from_numpy(yn)
    lg32, loss32, g32 = mlp_oracle.forward_backward(sd, x, y, 5, 1024)
    lg64, loss64, g64 = mlp_oracle.forward_backward(sd, x, y, 5, 1024, dtype=torch.float64)
    model = pkg.Model(hidden_layers=5, features=1024); model.load_state_dict(sd); model = model.cuda()
    loss = model.loss_and_grads(x.cuda(), y.cuda())
    assert abs(loss.item() - loss64.item()) <= 1e-5 * abs(loss64.item())
    with torch.no_grad():
        logits = model(x.cuda())
    assert (logits.cpu().double() - lg64).abs().max().item() <= 1e-5 * lg64.abs().max().item() + 2 * (lg32.double() - lg64).abs().max().item()
    top2 = lg64.topk(2, 1).values
    decided = (top2[:, 0] - top2[:, 1]) > 1e-5 * lg64.abs().max()
    assert torch.equal(logits.argmax(1).cpu()[decided], lg64.argmax(1)[decided]), "argmax must match wherever fp64 is decided"
    # (a) our masks, rebuilt from the public building blocks: the fused epilogue computes relu(acc + bias) from the same GEMM
    layers = ["input_layer"] + [f"hidden_layers.{i}" for i in range(5)]
    h = x.cuda().reshape(B, -1)
    masks = []
    for name in layers:
        a_hi, a_lo = native.split_tf32(h)
        b_hi, b_lo = native.split_tf32(sd[name + ".weight"].cuda())
        h = torch.relu(native.gemm_3xtf32(a_hi, a_lo, b_hi, b_lo) + sd[name + ".bias"].cuda())
        hh, hl = native.split_tf32(h)
        h = hh + hl                                   # activations travel as tf32 hi/lo planes
        masks.append((h > 0).double())
    p64 = {k: v.double().cuda().requires_grad_(True) for k, v in sd.items()}
    h64 = x.cuda().double().reshape(B, -1)
    for name, m in zip(layers, masks):
        h64 = (h64 @ p64[name + ".weight"].t() + p64[name + ".bias"]) * m
    loss_m = torch.nn.functional.cross_entropy(h64 @ p64["final_layer.weight"].t() + p64["final_layer.bias"], y.cuda())
    loss_m.backward()
    for k, p in model.named_parameters():
        gm = p64[k].grad
        err = (p.grad.double() - gm).abs().max().item()
        assert err <= 1e-5 * gm.abs().max().item(), f"{k}: {err:.3e} vs scale {gm.abs().max().item():.3e} (same-mask fp64 backward)"
    # (b) plain oracle, flip allowance
    for k, p in model.named_parameters():
        scale = g64[k].abs().max().item()
        err = (p.grad.double().cpu() - g64[k]).abs().max().item()
        own = (g32[k].double() - g64[k]).abs().max().item()
        assert err <= 1e-5 * scale + 2 * own + 4 * scale / B, f"{k}: err {err:.3e}, scale {scale:.3e}"
    # bit-identical reruns (no atomics anywhere)
    g1 = model.flat_grads.clone()
    model.loss_and_grads(x.cuda(), y.cuda())
    assert torch.equal(g1, model.flat_grads)


@pytest.mark.parametrize("B", [30, 1, 7])
def test_mlp_ragged_batch_is_padded_exactly(native, B):
    """The reference DataLoader's last batch is whatever is left (drop_last=False): model(x) must take any size."""
    import pytorch_distributed_examples_b200 as pkg
    from oracle import mlp_oracle, synth_oracle
    torch.manual_seed(2)
    sd = {k: v.detach().clone() for k, v in mlp_oracle.OracleMLP(1, 128).state_dict().items()}
    xn, yn = synth_oracle.synth_batch(5, np.arange(B))
    x, y = torch.from_numpy(xn), torch.from_numpy(yn)
    lg32, loss32, g32 = mlp_oracle.forward_backward(sd, x, y, 1, 128)
    _, _, g64 = mlp_oracle.forward_backward(sd, x, y, 1, 128, dtype=torch.float64)
    model = pkg.Model(1, 128); model.load_state_dict(sd); model = model.cuda()
    logits = model(x.cuda())
    assert logits.shape == (B, 10)
    loss = torch.nn.functional.cross_entropy(logits, y.cuda())
    loss.backward()
    assert abs(loss.item() - loss32.item()) <= 1e-5 * abs(loss32.item())
    for k, p in model.named_parameters():
        scale = g64[k].abs().max().item()
        assert (p.grad.double().cpu() - g64[k]).abs().max().item() <= 1e-5 * scale + 2 * (g32[k].double() - g64[k]).abs().max().item() + 4 * scale / max(B, 64), k
    with torch.no_grad():
        assert torch.equal(model(x.cuda()), logits.detach())


def test_mlp_two_outstanding_forwards_each_keep_their_activations(native):
    """Two forwards before one backward (a second micro-batch, or an evaluation pass inside the training step): every
    autograd forward owns its activation workspace, so the summed loss back-propagates through both correctly."""
    import pytorch_distributed_examples_b200 as pkg
    from oracle import mlp_oracle, synth_oracle
    torch.manual_seed(3)
    ref = mlp_oracle.OracleMLP(2, 128)
    sd = {k: v.detach().clone() for k, v in ref.state_dict().items()}
    xn, yn = synth_oracle.synth_batch(9, np.arange(96))
    x, y = torch.from_numpy(xn), torch.from_numpy(yn)
    xa, xb, ya, yb = x[:64], x[64:], y[:64], y[64:]
    ref = ref.double()
    (torch.nn.functional.cross_entropy(ref(xa.double()), ya) + 0.5 * torch.nn.functional.cross_entropy(ref(xb.double()), yb)).backward()
    model = pkg.Model(2, 128); model.load_state_dict(sd); model = model.cuda()
    la, lb = model(xa.cuda()), model(xb.cuda())                      # second forward BEFORE the first one's backward
    with torch.no_grad():
        model(xb.cuda())                                             # and an inference pass in between
    (torch.nn.functional.cross_entropy(la, ya.cuda()) + 0.5 * torch.nn.functional.cross_entropy(lb, yb.cuda())).backward()
    for (k, p), (_, q) in zip(model.named_parameters(), ref.named_parameters()):
        scale = q.grad.abs().max().item()
        assert (p.grad.double().cpu() - q.grad).abs().max().item() <= 1e-5 * scale + 4 * scale / 64, k


@pytest.mark.parametrize("H,Fe,B", [(1, 100, 36), (2, 36, 64), (0, 260, 132)])
def test_mlp_odd_shapes_against_oracle(native, H, Fe, B):
    """Feature counts that are multiples of 4 but not of the 32 / 128 tile sizes, and batches that are not multiples of
    32: the TMA units clip / zero-fill the ragged tiles (K, M and N edges of all three GEMM layouts)."""
    import pytorch_distributed_examples_b200 as pkg
    from oracle import mlp_oracle, synth_oracle
    torch.manual_seed(H * 1000 + Fe)
    sd = {k: v.detach().clone() for k, v in mlp_oracle.OracleMLP(H, Fe).state_dict().items()}
    xn, yn = synth_oracle.synth_batch(11, np.arange(B))
    x, y = torch.from_numpy(xn), torch.from_numpy(yn)
    lg32, loss32, g32 = mlp_oracle.forward_backward(sd, x, y, H, Fe)
    lg64, loss64, g64 = mlp_oracle.forward_backward(sd, x, y, H, Fe, dtype=torch.float64)
    model = pkg.Model(hidden_layers=H, features=Fe); model.load_state_dict(sd); model = model.cuda()
    loss = model.loss_and_grads(x.cuda(), y.cuda())
    assert abs(loss.item() - loss64.item()) <= 1e-5 * abs(loss64.item())
    for k, p in model.named_parameters():
        scale = g64[k].abs().max().item()
        err = (p.grad.double().cpu() - g64[k]).abs().max().item()
        assert err <= 1e-5 * scale + 2 * (g32[k].double() - g64[k]).abs().max().item() + 4 * scale / max(B, 64), (k, err, scale)
    with torch.no_grad():
        logits = model(x.cuda())
    assert (logits.cpu().double() - lg64).abs().max().item() <= 1e-5 * lg64.abs().max().item() + 2 * (lg32.double() - lg64).abs().max().item()
