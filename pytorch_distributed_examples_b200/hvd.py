"""``hvd`` — the slice of the ``horovod.torch`` API the reference scripts call, re-built B200-native.

Call sites replaced (ref: horovod/mnist_horovod.py, horovod/horovod_mnist_elastic.py):
    hvd.init()                                             :28 / :35
    hvd.size(), hvd.rank(), hvd.local_rank()               :42 / :41,51,73
    hvd.DistributedOptimizer(opt, named_parameters=...)    :53 / :42
    hvd.broadcast_parameters(model.state_dict(), 0)        :56
    hvd.elastic.run / hvd.elastic.TorchState / commit ...  elastic :55,72,104-105   (see elastic.py)
Semantics restated from Horovod's public API (the package itself is not installable here — SURVEY §8(c)):
``DistributedOptimizer.step()`` == average every ``p.grad`` over all ranks, then the wrapped optimizer's
update; ``broadcast_parameters`` == every rank ends with rank ``root_rank``'s tensors.

B200 design: gradients are gathered into ONE flat bucket (zero-copy for our ``Net``, whose backward writes
them there; one flatten kernel otherwise), ONE ncclAllReduce(sum) over NVLink, then ONE fused kernel that
applies 1/N and the optimizer rule in place (optim.py / csrc/update.cu).
"""
from __future__ import annotations

import torch

from . import _native, comm as _comm
from .bucket import BucketTable, FlatLayout, arena_of, flatten_into, unflatten_from
from .optim import FusedUpdater, _arena_from

Average, Sum = "Average", "Sum"


# A collective failed, reported an asynchronous error or missed its deadline (peer lost): raised by every collective /
# peer-memory entry point of the C-ABI wrapper and by Communicator.wait().  elastic.run restores the last commit and
# re-forms the ring over the survivors (ref: horovod/horovod_mnist_elastic.py:55).
HorovodInternalError = _native.CollectiveError


class HostsUpdatedInterrupt(RuntimeError):
    """Membership change was announced; raised from state.commit() on every rank at the same commit."""

    def __init__(self, skip_sync=False):
        super().__init__("hosts updated")
        self.skip_sync = skip_sync


class _Context:
    def __init__(self):
        self.initialized = False
        self.job_world = 1      # processes in the job (fixed)
        self.global_rank = 0
        self.local_rank = 0
        self.store = None
        self.comm = None        # current Communicator (None while this process is a standby)
        self.generation = 0     # membership-plan generation this ring was formed for
        self.attempt = 0        # re-formations within one plan generation (fresh ncclUniqueId each)


_ctx = _Context()


def init(ranks=None):
    """ref: horovod/mnist_horovod.py:28.  Reads the launcher's environment, binds this process to its GPU
    and forms the NCCL communicator.  ``ranks``: optional subset of global ranks forming the first ring."""
    if _ctx.initialized:
        return
    world, rank, local = _comm.env_world()
    _ctx.job_world, _ctx.global_rank, _ctx.local_rank = world, rank, local
    if torch.cuda.is_available():
        torch.cuda.set_device(local % torch.cuda.device_count())
        _native.get().bind_device(torch.cuda.current_device())
    _ctx.store = _comm.default_store(world)
    _ctx.generation, _ctx.attempt = 0, 0
    members = list(range(world)) if ranks is None else list(ranks)
    _ctx.comm = _comm.Communicator(_ctx.store, rank, members, "0.0") if rank in members else None
    _ctx.initialized = True


def shutdown():
    if _ctx.comm is not None:
        _ctx.comm.destroy()
    _ctx.comm = None
    _ctx.initialized = False


def is_initialized():
    return _ctx.initialized


def _require():
    if not _ctx.initialized:
        raise ValueError("Horovod has not been initialized; use hvd.init().")
    if _ctx.comm is None:
        raise HorovodInternalError("this process is not a member of the current ring")
    return _ctx.comm


def size():
    return _require().world


def rank():
    return _require().rank


def local_rank():
    if not _ctx.initialized:
        raise ValueError("Horovod has not been initialized; use hvd.init().")
    return _ctx.local_rank


def local_size():
    return torch.cuda.device_count() if torch.cuda.is_available() else 1


def cross_rank():
    return 0


def reform(ranks, generation=None):
    """Re-form the ring over ``ranks`` (global ranks).  Processes not listed become standbys (comm None).
    This is what hvd.shutdown()+hvd.init() do inside horovod's elastic reset."""
    if _ctx.comm is not None:
        if _ctx.comm.broken is None:
            torch.cuda.synchronize() if torch.cuda.is_available() else None
        else:                                 # the ring was aborted: its kernels have exited, nothing orderly to wait for
            try:
                torch.cuda.synchronize() if torch.cuda.is_available() else None
            except RuntimeError:
                pass
        _ctx.comm.destroy()
    generation = _ctx.generation if generation is None else generation
    _ctx.attempt = _ctx.attempt + 1 if generation == _ctx.generation else 0
    _ctx.generation = generation
    ranks = list(ranks)
    tag = f"{_ctx.generation}.{_ctx.attempt}"
    _ctx.comm = (_comm.Communicator(_ctx.store, _ctx.global_rank, ranks, tag)
                 if _ctx.global_rank in ranks else None)


# -- tensor collectives ------------------------------------------------------------------------------
def allreduce_(tensor, average=None, name=None, op=None):
    c = _require()
    avg = (op in (None, Average)) if average is None else bool(average)
    c.allreduce_sum_(tensor)
    if avg and c.world > 1:
        if tensor.dtype in (torch.float32, torch.bfloat16):
            _native.get().scale(tensor, 1.0 / c.world)
        else:
            tensor.div_(c.world, rounding_mode="floor")
    return tensor


def allreduce(tensor, average=None, name=None, op=None):
    return allreduce_(tensor.clone(), average, name, op)


def broadcast_(tensor, root_rank, name=None):
    return _require().broadcast_(tensor, root_rank)


def broadcast(tensor, root_rank, name=None):
    return broadcast_(tensor.clone(), root_rank, name)


def broadcast_object(obj, root_rank=0, name=None):
    return _require().broadcast_object(obj, root_rank)


def barrier():
    c = _require()
    if c.world > 1:
        t = torch.zeros(1, dtype=torch.float32, device="cuda" if torch.cuda.is_available() else "cpu")
        c.allreduce_sum_(t)
        if t.is_cuda:
            torch.cuda.current_stream().synchronize()


def _broadcast_tensors(tensors, root_rank):
    """One collective for a whole list: zero-copy if the tensors alias one arena, else flatten/unflatten."""
    c = _require()
    tensors = [t for t in tensors if isinstance(t, torch.Tensor)]
    if c.world <= 1 or not tensors:
        return
    by_dtype = {}
    for t in tensors:
        by_dtype.setdefault(t.dtype, []).append(t)
    for dtype, ts in by_dtype.items():
        layout = FlatLayout.for_tensors(ts)
        t0 = ts[0]
        if (arena_of(ts, layout) is not None and t0.storage_offset() >= layout.offsets[0] and
                t0.untyped_storage().nbytes() >= (t0.storage_offset() + layout.total) * t0.element_size()):
            c.broadcast_(_arena_from(t0, layout.offsets[0], layout.total), root_rank)
        elif len(ts) == 1:
            if t0.is_contiguous():
                c.broadcast_(t0, root_rank)
            else:
                tmp = t0.contiguous(); c.broadcast_(tmp, root_rank); t0.copy_(tmp)
        else:
            ts = [t if t.is_contiguous() else t.contiguous() for t in ts]
            flat = torch.zeros(layout.total, dtype=dtype, device=t0.device)
            table = BucketTable(t0.device)
            flatten_into(flat, ts, layout, table)
            c.broadcast_(flat, root_rank)
            unflatten_from(flat, ts, layout, table)


def broadcast_parameters(params, root_rank):
    """ref: horovod/mnist_horovod.py:56.  ``params``: ``model.state_dict()`` or ``model.named_parameters()``."""
    if isinstance(params, dict):
        items = list(params.values())
    else:
        items = [p for _, p in params]
    _broadcast_tensors([p.data if isinstance(p, torch.nn.Parameter) else p for p in items], root_rank)


def broadcast_optimizer_state(optimizer, root_rank):
    """Every rank ends with root's optimizer state (arenas + step) and param_group hyper-parameters."""
    c = _require()
    fused = getattr(optimizer, "_fused", None)
    if fused is None:
        raise TypeError("broadcast_optimizer_state expects an hvd.DistributedOptimizer / FusedOptimizer")
    hyper = [{k: v for k, v in g.items() if k != "params"} for g in optimizer.param_groups]
    hyper = c.broadcast_object(hyper, root_rank)
    for g, h in zip(optimizer.param_groups, hyper):
        g.update(h)
    state_arenas = []
    for g in fused.groups:
        g.ensure_state()
        state_arenas += [a for a in (g.state1, g.state2, g.step_dev) if a is not None]
    for a in state_arenas:
        c.broadcast_(a, root_rank)
    fused.set_steps_from_device()


# -- DistributedOptimizer -----------------------------------------------------------------------------
class _DistributedOptimizerMixin:
    """Methods grafted onto a dynamic subclass of the wrapped optimizer's class (like horovod does), so
    ``isinstance(opt, torch.optim.SGD)`` stays true and ``param_groups`` is the wrapped object's own list."""

    def _ddp_setup(self):
        self._fused = FusedUpdater(self)
        self._synchronized = False
        self._use_step_dev = False

    def synchronize(self):
        """Gather gradients into the bucket(s), allreduce(sum), average, make p.grad show the result."""
        c = _require()
        for g in self._fused.groups:
            bucket = g.gather()
            if c.world > 1:
                c.allreduce_sum_(bucket)
                _native.get().scale(bucket, 1.0 / c.world)
            g.scatter()
        self._synchronized = True

    def step(self, closure=None):
        """ref: horovod/mnist_horovod.py:64 — allreduce(Average) of every gradient, then the update; here:
        bucket allreduce(sum) + ONE fused (1/N)*g+update kernel per param group."""
        loss = closure() if closure is not None else None
        c = _require()
        if self._synchronized:           # user called synchronize() first: bucket already averaged
            for g in self._fused.groups:
                g.update(1.0, self._use_step_dev)
        else:
            for g in self._fused.groups:
                bucket = g.gather()
                if c.world > 1:
                    c.allreduce_sum_(bucket)
                g.update(1.0 / c.world, self._use_step_dev)
        self._synchronized = False
        return loss

    def zero_grad(self, set_to_none: bool = True):
        """ref: horovod/mnist_horovod.py:60 — free: backward overwrites the bucket, nothing to clear."""
        for group in self.param_groups:
            for p in group["params"]:
                p.grad = None

    def state_dict(self):
        self._fused.export_state()
        return super().state_dict()

    def load_state_dict(self, sd):
        super().load_state_dict(sd)
        self._fused.import_state()


def DistributedOptimizer(optimizer, named_parameters=None, compression=None, backward_passes_per_step=1,
                         op=Average, gradient_predivide_factor=1.0, **_ignored):
    """ref: horovod/mnist_horovod.py:53 ; horovod_mnist_elastic.py:42."""
    if backward_passes_per_step != 1:
        raise NotImplementedError("backward_passes_per_step != 1")
    if op not in (Average, None):
        raise NotImplementedError("only op=Average (the reference's default) is implemented")
    base = optimizer.__class__
    cls = type(base.__name__, (_DistributedOptimizerMixin, base), {})
    obj = cls.__new__(cls)
    obj.__dict__.update(optimizer.__dict__)     # shares param_groups / state / defaults with the wrapped object
    obj._ddp_setup()
    return obj


from . import elastic  # noqa: E402  (hvd.elastic.*)
