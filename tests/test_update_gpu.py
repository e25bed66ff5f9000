"""Fused (1/N)*g + optimizer update vs torch.optim (the reference's optimizers), through the C-ABI."""
import pytest
import torch

from helpers import assert_close_scaled
from oracle import net_oracle
from pytorch_distributed_examples_b200 import _native

pytestmark = pytest.mark.gpu

CASES = [
    ("sgd", _native.OPT_SGD, dict(lr=0.01)),                                   # ref: mnist_horovod.py:50
    ("sgd", _native.OPT_SGD, dict(lr=0.01, momentum=0.9)),
    ("sgd", _native.OPT_SGD, dict(lr=0.05, momentum=0.9, nesterov=True, weight_decay=1e-3)),
    ("sgd", _native.OPT_SGD, dict(lr=0.05, momentum=0.8, dampening=0.1)),
    ("adam", _native.OPT_ADAM, dict(lr=1e-3)),                                 # ref: mnist_ddp_elastic.py:173
    ("adam", _native.OPT_ADAM, dict(lr=1e-3, weight_decay=1e-2)),
    ("adamw", _native.OPT_ADAMW, dict(lr=0.01 / 2 ** 0.5)),                    # ref: horovod_mnist_elastic.py:41
    ("adadelta", _native.OPT_ADADELTA, dict(lr=1.0)),
    ("adadelta", _native.OPT_ADADELTA, dict(lr=0.5, rho=0.95, weight_decay=1e-3)),
]


def _hyper(kind, kw):
    h = _native.Hyper()
    h.lr = kw["lr"]; h.weight_decay = kw.get("weight_decay", 0.01 if kind == "adamw" else 0.0)
    h.momentum = kw.get("momentum", 0.0); h.dampening = kw.get("dampening", 0.0); h.nesterov = int(kw.get("nesterov", False))
    b = kw.get("betas", (0.9, 0.999)); h.beta1, h.beta2 = b
    h.eps = kw.get("eps", 1e-6 if kind == "adadelta" else 1e-8); h.rho = kw.get("rho", 0.9)
    return h


@pytest.mark.parametrize("kind,opt,kw", CASES)
@pytest.mark.parametrize("n,world", [(21848, 1), (1003, 2), (262147, 8)])
def test_fused_update_matches_torch_optim(native, kind, opt, kw, n, world):
    torch.manual_seed(n)
    p_ref = torch.nn.Parameter(torch.randn(n))
    ref = net_oracle.make_optimizer(kind, [p_ref], **kw)
    p = p_ref.detach().clone().cuda()
    s1 = torch.zeros(n, device="cuda"); s2 = torch.zeros(n, device="cuda")
    h = _hyper(kind, kw)
    for step in range(1, 6):
        gsum = torch.randn(n) * world                  # what the allreduce(sum) leaves in the bucket
        p_ref.grad = gsum / world
        ref.step()
        native.fused_update(opt, gsum.cuda(), p, s1, s2, n, 1.0 / world, h, step=step)
    torch.cuda.synchronize()
    assert_close_scaled(p, p_ref.detach(), rtol=1e-5 if kind != "adadelta" else 3e-5, name=f"{kind} {kw}")


def test_step_from_device_counter_and_bf16_grads(native):
    n = 4096
    torch.manual_seed(0)
    p_ref = torch.nn.Parameter(torch.randn(n))
    ref = torch.optim.AdamW([p_ref], lr=1e-2, foreach=False)
    p = p_ref.detach().clone().cuda()
    s1 = torch.zeros(n, device="cuda"); s2 = torch.zeros(n, device="cuda")
    step_dev = torch.zeros(1, dtype=torch.int64, device="cuda")
    h = _hyper("adamw", dict(lr=1e-2))
    for _ in range(4):
        g = torch.randn(n).bfloat16()
        p_ref.grad = g.float()
        ref.step()
        native.fused_update(_native.OPT_ADAMW, g.cuda(), p, s1, s2, n, 1.0, h, step=0, step_dev=step_dev)
        native.step_increment(step_dev)
    assert step_dev.item() == 4
    assert_close_scaled(p, p_ref.detach(), name="adamw bf16 grads / device step")


def test_write_back_and_scale(native):
    n = 1001
    g = torch.randn(n, device="cuda"); g0 = g.clone()
    p = torch.zeros(n, device="cuda")
    h = _hyper("sgd", dict(lr=1.0)); h.write_back_grad = 1
    native.fused_update(_native.OPT_SGD, g, p, None, None, n, 0.25, h, step=1)
    assert torch.equal(g, g0 * 0.25) and torch.equal(p, -(g0 * 0.25))
    b = torch.randn(777, device="cuda"); b0 = b.clone()
    native.scale(b, 0.125)
    assert torch.equal(b, b0 * 0.125)
    with pytest.raises(RuntimeError):
        native.fused_update(_native.OPT_ADAM, g, p, None, None, n, 1.0, h, step=1)   # missing state -> EINVAL


def test_fused_optimizer_wrapper_on_foreign_module(native):
    """The MLP of ref: pytorch_elastic/mnist_ddp_elastic.py:133-159 (small widths) through FusedOptimizer: params
    are re-homed into one arena, grads flattened by the bucket kernel, result equals torch.optim.Adam."""
    import copy
    import pytorch_distributed_examples_b200 as pkg
    torch.manual_seed(3)
    mlp = torch.nn.Sequential(torch.nn.Linear(784, 33), torch.nn.ReLU(), torch.nn.Linear(33, 17), torch.nn.ReLU(),
                              torch.nn.Linear(17, 10))
    ref = copy.deepcopy(mlp)
    ropt = torch.optim.Adam(ref.parameters(), lr=1e-3, foreach=False)
    mlp = mlp.cuda()
    fopt = pkg.FusedOptimizer(torch.optim.Adam(mlp.parameters(), lr=1e-3))
    crit = torch.nn.CrossEntropyLoss()
    for k in range(3):
        x = torch.randn(16, 784); y = torch.randint(0, 10, (16,))
        ropt.zero_grad(); crit(ref(x), y).backward(); ropt.step()
        fopt.zero_grad(); crit(mlp(x.cuda()), y.cuda()).backward(); fopt.step()
    for a, b in zip(mlp.parameters(), ref.parameters()):
        # gradients here come from cuBLAS vs MKL; Adam amplifies their last-bit differences (see test_train_gpu)
        err = (a.detach().cpu() - b.detach()).abs().max().item()
        assert err <= 1e-5 * b.abs().max().item() + 1e-2 * 1e-3, err
    sd = fopt.state_dict()
    assert "exp_avg" in sd["state"][0] and float(sd["state"][0]["step"]) == 3.0
