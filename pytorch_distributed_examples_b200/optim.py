"""Fused "average + optimizer update" over flat arenas, driven by an ordinary ``torch.optim`` object.

The reference scripts build stock optimizers — ``optim.SGD(lr=0.01)`` (ref: horovod/mnist_horovod.py:50),
``optim.AdamW(lr=0.01/sqrt(size))`` (ref: horovod_mnist_elastic.py:41), ``optim.Adam(lr=1e-3)``
(ref: pytorch_elastic/mnist_ddp_elastic.py:173) — and call ``optimizer.step()``.  ``FusedUpdater`` keeps that
object as the source of truth for hyper-parameters (``param_groups`` stay live, so the elastic
``on_state_reset`` LR rescale, ref: horovod_mnist_elastic.py:80-82, just works) but replaces the per-tensor
Python loop with ONE ddp_fused_update launch per param group that reads the (allreduced) gradient bucket
once and updates parameters and optimizer state in place (csrc/update.cu).
"""
from __future__ import annotations

from typing import List

import torch

from . import _native
from .bucket import BucketTable, FlatLayout, arena_of, flatten_into, unflatten_from


def _arena_from(t0: torch.Tensor, off0: int, total: int) -> torch.Tensor:
    """A 1-D tensor over the storage ``t0`` lives in, starting ``off0`` elements before it."""
    return torch.empty(0, dtype=t0.dtype, device=t0.device).set_(
        t0.untyped_storage(), t0.storage_offset() - off0, (total,), (1,))


def optimizer_kind(opt: torch.optim.Optimizer) -> int:
    if isinstance(opt, torch.optim.AdamW):
        return _native.OPT_ADAMW
    if isinstance(opt, torch.optim.Adam):
        return _native.OPT_ADAM
    if isinstance(opt, torch.optim.SGD):
        return _native.OPT_SGD
    if isinstance(opt, torch.optim.Adadelta):
        return _native.OPT_ADADELTA
    raise TypeError(f"no fused sm_100a update for {type(opt).__name__} "
                    "(supported: SGD, Adam, AdamW, Adadelta)")


class _Group:
    """One param group laid out flat: params arena, gradient bucket, up to two state arenas."""

    def __init__(self, kind: int, group: dict):
        self.group = group
        self.kind = kind
        self.params: List[torch.nn.Parameter] = [p for p in group["params"] if p.requires_grad]
        if not self.params:
            self.layout = None
            return
        p0 = self.params[0]
        if not p0.is_cuda and not getattr(_native.get(), "allows_cpu", False):   # only a test double allows CPU
            raise RuntimeError("fused optimizer update needs CUDA parameters (no CPU fallback)")
        if any(p.dtype != torch.float32 for p in self.params):
            raise RuntimeError("fused optimizer update keeps fp32 master parameters")
        self.layout = FlatLayout.for_tensors(self.params)
        datas = [p.data for p in self.params]
        if arena_of(datas, self.layout) is not None and self._storage_covers(datas[0], self.layout):
            self.param_arena = _arena_from(datas[0], self.layout.offsets[0], self.layout.total)
        else:
            self.param_arena = torch.zeros(self.layout.total, dtype=torch.float32, device=p0.device)
            with torch.no_grad():
                for p, v in zip(self.params, self.layout.views(self.param_arena)):
                    v.copy_(p.data)
                    p.data = v
        self.bucket = None           # gradient bucket (adopted from the grads or owned)
        self.adopted = False
        self.own_bucket = None
        self.table = BucketTable(p0.device)
        self.state1 = self.state2 = None
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=p0.device)
        self.steps = 0
        self.graph_mode = False      # True: the device counter is the truth (CUDA-graph replays)

    @staticmethod
    def _storage_covers(t0, layout):
        need = (t0.storage_offset() - layout.offsets[0] + layout.total) * t0.element_size()
        return t0.storage_offset() >= layout.offsets[0] and t0.untyped_storage().nbytes() >= need

    def needs_state(self):
        if self.kind == _native.OPT_SGD:
            return (self.group.get("momentum", 0) != 0, False)
        return (True, True)

    def ensure_state(self):
        n1, n2 = self.needs_state()
        if n1 and self.state1 is None:
            self.state1 = torch.zeros_like(self.param_arena)
        if n2 and self.state2 is None:
            self.state2 = torch.zeros_like(self.param_arena)

    def hyper(self, write_back=False) -> _native.Hyper:
        g = self.group
        h = _native.Hyper()
        h.lr = float(g["lr"])
        h.weight_decay = float(g.get("weight_decay", 0.0))
        h.write_back_grad = int(write_back)
        if self.kind == _native.OPT_SGD:
            h.momentum = float(g.get("momentum", 0.0)); h.dampening = float(g.get("dampening", 0.0))
            h.nesterov = int(bool(g.get("nesterov", False)))
            if g.get("maximize", False):
                raise RuntimeError("maximize=True is not supported by the fused update")
        elif self.kind in (_native.OPT_ADAM, _native.OPT_ADAMW):
            h.beta1, h.beta2 = (float(b) for b in g["betas"]); h.eps = float(g["eps"])
            if g.get("amsgrad", False) or g.get("maximize", False):
                raise RuntimeError("amsgrad / maximize are not supported by the fused update")
        else:
            h.rho = float(g["rho"]); h.eps = float(g["eps"])
        return h

    # -- gradients -> bucket --------------------------------------------------------------------------
    def gather(self, grad_dtype=torch.float32) -> torch.Tensor:
        """Make ``self.bucket`` hold this step's gradients: zero-copy when they were born in one arena
        (our Net), else one flatten kernel launch."""
        grads = [p.grad for p in self.params]
        if any(g is None for g in grads):
            missing = [i for i, g in enumerate(grads) if g is None]
            raise RuntimeError(f"parameters {missing} have no gradient; every parameter must take part in backward")
        if arena_of(grads, self.layout) is not None and self._storage_covers(grads[0], self.layout):
            self.bucket = _arena_from(grads[0], self.layout.offsets[0], self.layout.total)
            self.adopted = True
        else:
            if self.own_bucket is None or self.own_bucket.dtype != grads[0].dtype:
                self.own_bucket = torch.zeros(self.layout.total, dtype=grads[0].dtype, device=grads[0].device)
            flatten_into(self.own_bucket, grads, self.layout, self.table)
            self.bucket = self.own_bucket
            self.adopted = False
        return self.bucket

    def scatter(self):
        """bucket -> p.grad (after an explicit synchronize(), so users see the averaged gradients)."""
        if not self.adopted:
            unflatten_from(self.bucket, [p.grad for p in self.params], self.layout, self.table)

    def sync_step_to_device(self):
        """Make the device step counter agree with the host count (host is the truth outside graph mode)."""
        if not self.graph_mode:
            self.step_dev.fill_(self.steps)

    def update(self, inv_world: float, use_step_dev=False):
        self.graph_mode = bool(use_step_dev)
        self.ensure_state()
        nat = _native.get()
        self.steps += 1
        hyper = self.hyper()
        hyper.bump_step_dev = int(bool(use_step_dev))     # the update kernel's last CTA advances the device counter
        nat.fused_update(self.kind, self.bucket, self.param_arena, self.state1, self.state2, self.layout.total,
                         inv_world, hyper, step=self.steps, step_dev=self.step_dev if use_step_dev else None)

    def update_peer(self, comm, use_step_dev=False):
        """allreduce(sum) over NVLink peer memory + 1/N + update in ONE kernel (csrc/peer.cu)."""
        self.graph_mode = bool(use_step_dev)
        self.ensure_state()
        nat = _native.get()
        self.steps += 1
        hyper = self.hyper()
        hyper.bump_step_dev = int(bool(use_step_dev))
        nat.peer_allreduce_update(comm.peer, self.kind, self.bucket, self.param_arena, self.state1, self.state2,
                                  self.layout.total, hyper, step=self.steps,
                                  step_dev=self.step_dev if use_step_dev else None)

    # -- torch.optim state interop (state_dict / load_state_dict / elastic commit) --------------------
    _STATE_KEYS = {_native.OPT_SGD: ("momentum_buffer", None), _native.OPT_ADAM: ("exp_avg", "exp_avg_sq"),
                   _native.OPT_ADAMW: ("exp_avg", "exp_avg_sq"), _native.OPT_ADADELTA: ("square_avg", "acc_delta")}

    def export_state(self, state: dict):
        """Publish arena views as the wrapped optimizer's ``state[p]`` so ``state_dict()`` has the usual keys."""
        if self.layout is None or self.steps == 0:
            return
        k1, k2 = self._STATE_KEYS[self.kind]
        v1 = self.layout.views(self.state1) if self.state1 is not None else None
        v2 = self.layout.views(self.state2) if self.state2 is not None else None
        for i, p in enumerate(self.params):
            st = state[p]
            if v1 is not None and k1:
                st[k1] = v1[i]
            if v2 is not None and k2:
                st[k2] = v2[i]
            if self.kind != _native.OPT_SGD:
                st["step"] = torch.tensor(float(self.steps))

    def import_state(self, state: dict):
        """Adopt state loaded through ``load_state_dict`` back into the arenas."""
        if self.layout is None:
            return
        k1, k2 = self._STATE_KEYS[self.kind]
        steps = 0
        with torch.no_grad():
            for i, p in enumerate(self.params):
                st = state.get(p, {})
                if k1 and k1 in st and st[k1] is not None:
                    self.ensure_state()
                    self.layout.views(self.state1)[i].copy_(st[k1])
                    steps = max(steps, 1)
                if k2 and k2 in st:
                    self.ensure_state()
                    self.layout.views(self.state2)[i].copy_(st[k2])
                if "step" in st:
                    steps = max(steps, int(float(st["step"])))
        if steps:
            self.steps = steps
            self.step_dev.fill_(steps)


class FusedUpdater:
    """All param groups of one torch optimizer, flat."""

    def __init__(self, optimizer: torch.optim.Optimizer):
        self.optimizer = optimizer
        self.kind = optimizer_kind(optimizer)
        self.groups = [_Group(self.kind, g) for g in optimizer.param_groups]
        self.groups = [g for g in self.groups if g.layout is not None]

    def gather(self):
        return [g.gather() for g in self.groups]

    def scatter(self):
        for g in self.groups:
            g.scatter()

    def update(self, inv_world: float = 1.0, use_step_dev=False):
        for g in self.groups:
            g.update(inv_world, use_step_dev)

    def export_state(self):
        for g in self.groups:
            g.export_state(self.optimizer.state)

    def import_state(self):
        for g in self.groups:
            g.import_state(self.optimizer.state)

    def arenas(self):
        """Every persistent device arena (params, states, step) — what elastic commit/sync must cover."""
        out = []
        for g in self.groups:
            g.ensure_state()
            out.append(g.param_arena)
            if g.state1 is not None:
                out.append(g.state1)
            if g.state2 is not None:
                out.append(g.state2)
            out.append(g.step_dev)
        return out

    def set_steps_from_device(self):
        for g in self.groups:
            g.steps = int(g.step_dev.item())


class FusedOptimizer:
    """Single-process convenience: ``FusedOptimizer(torch.optim.SGD(model.parameters(), lr=0.01))`` gives an
    object with the optimizer's ``step()/zero_grad()/param_groups`` whose step is the fused kernel."""

    def __init__(self, optimizer: torch.optim.Optimizer):
        self.optimizer = optimizer
        self._fused = FusedUpdater(optimizer)

    @property
    def param_groups(self):
        return self.optimizer.param_groups

    def zero_grad(self, set_to_none: bool = True):
        self.optimizer.zero_grad(set_to_none=True)

    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self._fused.gather()
        self._fused.update(1.0)
        return loss

    def state_dict(self):
        self._fused.export_state()
        return self.optimizer.state_dict()

    def load_state_dict(self, sd):
        self.optimizer.load_state_dict(sd)
        self._fused.import_state()
