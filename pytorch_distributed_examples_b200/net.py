"""``Net`` — drop-in for the reference MNIST CNN (ref: horovod/mnist_horovod.py:9-25,
horovod_mnist_elastic.py:16-32) whose forward/backward run as hand-written sm_100a kernels.

Same constructor, same submodule attribute names (conv1, conv2, conv2_drop, fc1, fc2) and therefore the same
``state_dict()`` keys and the same default initialisation as the reference (the submodules ARE torch's
``nn.Conv2d``/``nn.Linear``, used only as parameter holders).  ``model(data)`` returns the [B,10]
log-probabilities wired into autograd, so the reference training loop
    output = model(data); loss = F.nll_loss(output, target); loss.backward(); optimizer.step()
(ref: horovod/mnist_horovod.py:60-64) runs unchanged.  On a CUDA device the eight parameters are views of
one flat fp32 arena (include/ddp_step.h layout) and the gradients are written by the backward kernels
straight into a flat gradient bucket — the buffer the allreduce and the fused optimizer update consume.
There is no CPU implementation: calling the module with CPU tensors raises.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _native

_PARAM_ORDER = ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias",
                "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")


class _NetFunction(torch.autograd.Function):
    """forward = K_F1..K_F3, backward = K_B3, K_B2d, K_B2w, K_B1, K_R  (csrc/net_fwd.cu, net_bwd.cu)."""

    @staticmethod
    def forward(ctx, x, net, mask2d, mask1d, *params):
        nat = _native.get()
        batch = x.shape[0]
        logp = torch.empty(batch, 10, dtype=torch.float32, device=x.device)
        ws = torch.empty(nat.workspace_bytes(batch), dtype=torch.uint8, device=x.device)
        training = net.training
        offset = net._dropout_calls
        if training:
            net._dropout_calls += 1
        nat.net_forward(net._arena, x, None, mask2d, mask1d, net.dropout_seed, offset, training, ws, logp)
        ctx.net = net
        ctx.training = training
        ctx.save_for_backward(x, logp, ws)
        return logp

    @staticmethod
    def backward(ctx, dlogp):
        nat = _native.get()
        x, logp, ws = ctx.saved_tensors
        net = ctx.net
        # Gradients are born in the flat bucket.  If a previous backward's gradients are still attached
        # (gradient accumulation), write to a temporary so autograd's in-place += stays correct.
        fresh = all(p.grad is None for p in net._param_list)
        bucket = net._grad_arena if fresh else torch.empty_like(net._grad_arena)
        nat.net_backward(net._arena, x, logp, dlogp.contiguous(), None, 1.0, ctx.training, ws,
                         net._get_scratch(x.shape[0]), bucket)
        grads = [bucket[o:o + n].view(p.shape) for p, o, n in zip(net._param_list, net._offsets, net._sizes)]
        return (None, None, None, None, *grads)


class Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 10, kernel_size=5)
        self.conv2 = nn.Conv2d(10, 20, kernel_size=5)
        self.conv2_drop = nn.Dropout2d()
        self.fc1 = nn.Linear(320, 50)
        self.fc2 = nn.Linear(50, 10)
        self.dropout_seed = torch.initial_seed() & (2 ** 63 - 1)
        self._dropout_calls = 0
        self._arena = None
        self._grad_arena = None
        self._scratch = None
        self._scratch_batch = 0
        self._offsets = self._sizes = None

    # -- flat arena management ----------------------------------------------------------------------
    @property
    def _param_list(self):
        return [self.conv1.weight, self.conv1.bias, self.conv2.weight, self.conv2.bias,
                self.fc1.weight, self.fc1.bias, self.fc2.weight, self.fc2.bias]

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        self._flatten()
        return out

    def _flatten(self):
        """(Re)build the flat parameter arena after the module moved to a CUDA device."""
        dev = self.conv1.weight.device
        if dev.type != "cuda":
            self._arena = self._grad_arena = self._scratch = None
            return
        nat = _native.get()
        self._offsets, self._sizes = nat.param_layout()
        arena = torch.zeros(_native.ARENA_ELEMS, dtype=torch.float32, device=dev)
        with torch.no_grad():
            for p, o, n in zip(self._param_list, self._offsets, self._sizes):
                arena[o:o + n].copy_(p.detach().reshape(-1).to(torch.float32))
                p.data = arena[o:o + n].view(p.shape)
                p.grad = None
        self._arena = arena
        self._grad_arena = torch.zeros_like(arena)
        self._scratch = None
        self._scratch_batch = 0

    def _check_arena(self):
        if self._arena is None:
            raise RuntimeError("Net runs only on a CUDA device (sm_100a kernels, no CPU fallback): call .cuda() first")
        base = self._arena.data_ptr()
        for p, o in zip(self._param_list, self._offsets):
            if p.data_ptr() != base + 4 * o:
                self._flatten()   # someone re-assigned .data: adopt the new values into a fresh arena
                break

    def _get_scratch(self, batch: int):
        if self._scratch is None or self._scratch_batch < batch:
            nat = _native.get()
            self._scratch = torch.zeros(nat.scratch_bytes(batch), dtype=torch.uint8, device=self._arena.device)
            self._scratch_batch = batch
        return self._scratch

    @property
    def flat_params(self) -> torch.Tensor:
        self._check_arena()
        return self._arena

    @property
    def flat_grads(self) -> torch.Tensor:
        self._check_arena()
        return self._grad_arena

    # -- the reference's forward --------------------------------------------------------------------
    def forward(self, x, mask2d=None, mask1d=None):
        """x [B,1,28,28] (or anything viewable to it).  mask2d [B,20] / mask1d [B,50] optionally inject the
        dropout keep-masks (parity tests); otherwise training mode draws them from Philox on the device."""
        self._check_arena()
        if not x.is_cuda:
            raise RuntimeError("Net.forward needs a CUDA tensor (no CPU fallback)")
        x = x.reshape(-1, 1, 28, 28).to(torch.float32).contiguous()
        if mask2d is not None:
            mask2d = mask2d.reshape(x.shape[0], 20).to(torch.float32).contiguous()
        if mask1d is not None:
            mask1d = mask1d.reshape(x.shape[0], 50).to(torch.float32).contiguous()
        return _NetFunction.apply(x, self, mask2d, mask1d, *self._param_list)

    # -- evaluation helper (ref: horovod_mnist_elastic.py:85-102 test(); mnist_ddp_elastic.py:117-130) ---
    @torch.no_grad()
    def count_correct(self, images, labels, correct_counter=None):
        """Forward in eval mode + fused argmax==label count; returns the device counter (int64[1])."""
        self._check_arena()
        nat = _native.get()
        images = images.reshape(-1, 1, 28, 28).to(torch.float32).contiguous()
        batch = images.shape[0]
        if correct_counter is None:
            correct_counter = torch.zeros(1, dtype=torch.int64, device=images.device)
        logp = torch.empty(batch, 10, dtype=torch.float32, device=images.device)
        ws = torch.empty(nat.workspace_bytes(batch), dtype=torch.uint8, device=images.device)
        nat.net_forward(self._arena, images, labels.contiguous(), None, None, 0, 0, False, ws, logp,
                        out_correct=correct_counter)
        return correct_counter
