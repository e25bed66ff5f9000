"""Torch-CPU restatement of the reference CNN ``Net`` with the dropout masks made explicit.

Follows ref: horovod/mnist_horovod.py:9-25 (identical in horovod_mnist_elastic.py:16-32):
    x = relu(max_pool2d(conv1(x), 2))                        :19
    x = relu(max_pool2d(conv2_drop(conv2(x)), 2))            :20   conv2_drop = nn.Dropout2d()  :14
    x = x.view(-1, 320)                                      :21
    x = relu(fc1(x))                                         :22
    x = dropout(x, training=self.training)                   :23
    x = fc2(x); return log_softmax(x)                        :24-25 (implicit dim -> 1 for 2-D input)
torch draws the Dropout2d noise as empty(B,20,1,1).bernoulli_(0.5)/0.5 and the F.dropout noise as
empty_like(x).bernoulli_(0.5)/0.5, in that order; ``draw_masks`` replays exactly those calls so a
seeded reference forward can be reproduced with injected masks (checked in gen_golden.py).
TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

PARAM_NAMES = ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias",
               "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")
PARAM_SHAPES = ((10, 1, 5, 5), (10,), (20, 10, 5, 5), (20,), (50, 320), (50,), (10, 50), (10,))


class OracleNet(nn.Module):
    """Same attribute names / state_dict keys as the reference Net (ref: horovod/mnist_horovod.py:10-16)."""

    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 10, kernel_size=5)
        self.conv2 = nn.Conv2d(10, 20, kernel_size=5)
        self.conv2_drop = nn.Dropout2d()
        self.fc1 = nn.Linear(320, 50)
        self.fc2 = nn.Linear(50, 10)

    def forward(self, x, mask2d=None, mask1d=None):
        """mask2d [B,20] / mask1d [B,50] in {0,1}: injected keep-masks (training mode).
        With masks None: training -> torch's own dropout RNG (reference behaviour), eval -> identity."""
        x = F.relu(F.max_pool2d(self.conv1(x), 2))
        y = self.conv2(x)
        if mask2d is not None:
            y = y * (mask2d.view(-1, 20, 1, 1) * 2.0)
        else:
            y = self.conv2_drop(y)
        x = F.relu(F.max_pool2d(y, 2))
        x = x.view(-1, 320)
        x = F.relu(self.fc1(x))
        if mask1d is not None:
            x = x * (mask1d * 2.0)
        else:
            x = F.dropout(x, training=self.training)
        x = self.fc2(x)
        return F.log_softmax(x, dim=1)


def draw_masks(batch: int, generator=None):
    """The two bernoulli draws Net.forward makes in training mode, in consumption order."""
    m2 = torch.empty(batch, 20, 1, 1).bernoulli_(0.5, generator=generator).view(batch, 20)
    m1 = torch.empty(batch, 50).bernoulli_(0.5, generator=generator)
    return m2, m1


def forward_backward(state_dict, x, target, mask2d=None, mask1d=None, training=True, dtype=torch.float32):
    """One forward + F.nll_loss + backward (ref: horovod/mnist_horovod.py:61-63).
    Returns (logp, loss, grads dict).  dtype=float64 gives the exact-arithmetic yardstick: at B=4096 torch's
    own fp32 CPU conv1 weight gradient (a 2.4M-term fp32 reduction) is ~1e-4 relative off it."""
    net = OracleNet().to(dtype)
    net.load_state_dict({k: v.to(dtype) for k, v in state_dict.items()})
    net.train(training)
    x = x.to(dtype)
    mask2d = None if mask2d is None else mask2d.to(dtype)
    mask1d = None if mask1d is None else mask1d.to(dtype)
    logp = net(x, mask2d, mask1d)
    loss = F.nll_loss(logp, target)
    loss.backward()
    grads = {k: p.grad.detach().clone() for k, p in net.named_parameters()}
    return logp.detach(), loss.detach(), grads


def make_optimizer(kind: str, params, **kw):
    """The optimizers the reference scripts construct (+ the two BASELINE.json names)."""
    if kind == "sgd":        # ref: horovod/mnist_horovod.py:50
        return torch.optim.SGD(params, lr=kw.get("lr", 0.01), momentum=kw.get("momentum", 0.0),
                               weight_decay=kw.get("weight_decay", 0.0), nesterov=kw.get("nesterov", False),
                               dampening=kw.get("dampening", 0.0), foreach=False)
    if kind == "adam":       # ref: pytorch_elastic/mnist_ddp_elastic.py:173
        return torch.optim.Adam(params, lr=kw.get("lr", 1e-3), betas=kw.get("betas", (0.9, 0.999)),
                                eps=kw.get("eps", 1e-8), weight_decay=kw.get("weight_decay", 0.0), foreach=False)
    if kind == "adamw":      # ref: horovod/horovod_mnist_elastic.py:41
        return torch.optim.AdamW(params, lr=kw.get("lr", 0.01), betas=kw.get("betas", (0.9, 0.999)),
                                 eps=kw.get("eps", 1e-8), weight_decay=kw.get("weight_decay", 0.01), foreach=False)
    if kind == "adadelta":   # named by BASELINE.json north_star; not in the reference
        return torch.optim.Adadelta(params, lr=kw.get("lr", 1.0), rho=kw.get("rho", 0.9),
                                    eps=kw.get("eps", 1e-6), weight_decay=kw.get("weight_decay", 0.0), foreach=False)
    raise ValueError(kind)


def train_steps(state_dict, xs, ys, masks2d, masks1d, opt_kind="sgd", world=1, **opt_kw):
    """k steps of the reference loop (ref: horovod/mnist_horovod.py:59-64) with injected masks.

    xs [k, W*B, 1,28,28]: with world>1 each step's batch is split into W equal shards, every shard's
    gradient is computed separately and the W gradients are AVERAGED before optimizer.step() — the
    semantics of hvd.DistributedOptimizer (ref :53) and of DDP's bucket allreduce (mnist_ddp_elastic.py:58).
    Returns (losses [k, W], final state_dict)."""
    net = OracleNet()
    net.load_state_dict(state_dict)
    net.train()
    opt = make_optimizer(opt_kind, net.parameters(), **opt_kw)
    losses = []
    for k in range(xs.shape[0]):
        xsh = xs[k].chunk(world)
        ysh = ys[k].chunk(world)
        m2 = masks2d[k].chunk(world) if masks2d is not None else [None] * world
        m1 = masks1d[k].chunk(world) if masks1d is not None else [None] * world
        acc = [torch.zeros_like(p) for p in net.parameters()]
        step_losses = []
        for r in range(world):
            opt.zero_grad()
            loss = F.nll_loss(net(xsh[r], m2[r], m1[r]), ysh[r])
            loss.backward()
            step_losses.append(loss.detach())
            for a, p in zip(acc, net.parameters()):
                a += p.grad
        for a, p in zip(acc, net.parameters()):
            p.grad = a / world
        opt.step()
        losses.append(torch.stack(step_losses))
    return torch.stack(losses), {k: v.detach().clone() for k, v in net.state_dict().items()}


def routing(state_dict, x, mask2d=None, mask1d=None, dtype=torch.float64):
    """Where the forward ROUTES gradients (ref: horovod/mnist_horovod.py:19-23): per 2x2 pool window the argmax
    (0..3 in window scan order) and the margin to the runner-up, per ReLU the pre-activation.  Two fp32
    implementations can only disagree on a route where the margin / |pre-activation| is within rounding noise,
    so parity tests count the disagreements and require every one of them to sit on such a near-tie."""
    net = OracleNet().to(dtype)
    net.load_state_dict({k: v.to(dtype) for k, v in state_dict.items()})
    x = x.to(dtype)

    def pool_route(y):                                   # y [B,C,H,W] -> idx [B,C,H/2,W/2] in 0..3, best, margin
        B, C, H, W = y.shape
        w = y.view(B, C, H // 2, 2, W // 2, 2).permute(0, 1, 2, 4, 3, 5).reshape(B, C, H // 2, W // 2, 4)
        srt, _ = w.sort(dim=-1, descending=True)
        best = srt[..., 0]
        idx = (w == best.unsqueeze(-1)).float().argmax(dim=-1)   # first maximum wins, like max_pool2d
        return idx, best, srt[..., 0] - srt[..., 1]

    with torch.no_grad():
        y1 = net.conv1(x)
        idx1, p1, m1 = pool_route(y1)
        a1 = F.relu(p1)
        y2 = net.conv2(a1)
        if mask2d is not None:
            y2 = y2 * (mask2d.to(dtype).view(-1, 20, 1, 1) * 2.0)
        idx2, p2, m2 = pool_route(y2)
        a2 = F.relu(p2).view(-1, 320)
        z1 = net.fc1(a2)
    return dict(idx1=idx1.reshape(x.shape[0], -1), pre1=p1.reshape(x.shape[0], -1), margin1=m1.reshape(x.shape[0], -1),
                idx2=idx2.reshape(x.shape[0], -1), pre2=p2.reshape(x.shape[0], -1), margin2=m2.reshape(x.shape[0], -1),
                pre_fc1=z1, scale1=y1.abs().max().item(), scale2=y2.abs().max().item(), scale_fc1=z1.abs().max().item())
