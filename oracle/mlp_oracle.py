"""CPU restatement of the DDP script's model and loop body (TEST INFRASTRUCTURE ONLY — see oracle/__init__.py).

ref: pytorch_elastic/mnist_ddp_elastic.py:133-159  `Model(hidden_layers, features)`:
     flatten -> relu(input_layer) -> relu(hidden_layers[i]) ... -> final_layer (logits, no softmax);
     :172-174 builds Model(hidden_layers=5, features=1024), Adam(lr=1e-3), nn.CrossEntropyLoss();
     :70-75 `_run_batch`: zero_grad, output = model(source), loss = criterion(output, targets), backward, step.
Pinned by oracle/gen_golden.py, which imports the reference module itself and asserts bit-for-bit equality.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import net_oracle


class OracleMLP(nn.Module):
    def __init__(self, hidden_layers=1, features=128, in_features=784, out_features=10):
        super().__init__()
        self.input_layer = nn.Linear(in_features, features)
        self.hidden_layers = nn.ModuleList([nn.Linear(features, features) for _ in range(hidden_layers)])
        self.final_layer = nn.Linear(features, out_features)

    def forward(self, x):
        h = F.relu(self.input_layer(x.reshape(x.size(0), -1)))
        for layer in self.hidden_layers:
            h = F.relu(layer(h))
        return self.final_layer(h)


def forward_backward(state_dict, x, target, hidden_layers, features, dtype=torch.float32):
    """logits, mean cross-entropy, gradients (ref :72-74).  dtype=float64 is the exact-arithmetic yardstick."""
    net = OracleMLP(hidden_layers, features, in_features=x[0].numel()).to(dtype)
    net.load_state_dict({k: v.to(dtype) for k, v in state_dict.items()})
    logits = net(x.to(dtype))
    loss = F.cross_entropy(logits, target)
    loss.backward()
    return logits.detach(), loss.detach(), {k: p.grad.detach().clone() for k, p in net.named_parameters()}


def train_steps(state_dict, xs, ys, hidden_layers, features, opt_kind="adam", world=1, **opt_kw):
    """k steps of `_run_batch` (ref :70-75); world > 1 averages the per-shard gradients like DDP does (:58).
    Returns (losses [k, W], final state_dict)."""
    net = OracleMLP(hidden_layers, features, in_features=xs[0][0].numel())
    net.load_state_dict(state_dict)
    opt = net_oracle.make_optimizer(opt_kind, net.parameters(), **opt_kw)
    losses = []
    for k in range(xs.shape[0]):
        acc = [torch.zeros_like(p) for p in net.parameters()]
        step_losses = []
        for xr, yr in zip(xs[k].chunk(world), ys[k].chunk(world)):
            opt.zero_grad()
            loss = F.cross_entropy(net(xr), yr)
            loss.backward()
            step_losses.append(loss.detach())
            for a, p in zip(acc, net.parameters()):
                a += p.grad
        for a, p in zip(acc, net.parameters()):
            p.grad = a / world
        opt.step()
        losses.append(torch.stack(step_losses))
    return torch.stack(losses), {k: v.detach().clone() for k, v in net.state_dict().items()}
