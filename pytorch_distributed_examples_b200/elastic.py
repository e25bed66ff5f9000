"""Elastic state: commit / restore / sync / reset — the ``hvd.elastic`` API of the reference.

Call sites (ref: horovod/horovod_mnist_elastic.py):
    @hvd.elastic.run                                 :55   retry loop around train(state)
    state.commit()                                   :72   every batches_per_commit batches
    state.batch / state.epoch                        :61-62,75-76  attribute access, assigned as loop variables
    hvd.elastic.TorchState(model, optimizer, batch=0, epoch=0)   :104
    state.register_reset_callbacks([on_state_reset]) :105  (LR <- lr/sqrt(new size), :80-82)
Restated Horovod semantics (package absent here):
    run:     loop { sync(); try fn(state) ; on HorovodInternalError: restore(); on HostsUpdatedInterrupt: keep;
                    reset ring; on_reset callbacks }
    commit:  save() an in-memory copy of model+optimizer state and the scalars, then check for host updates
    restore: roll back to the last commit;   sync: broadcast everything from rank 0
B200 design: model + optimizer state live in a few flat device arenas (optim.FusedUpdater.arenas()); commit is
an async device->pinned-host copy of those arenas (ddp_state_snapshot) fenced by a CUDA event, restore the
reverse copy, sync one ncclBroadcast per arena.  Membership changes on a single 8-GPU host are announced
through the rendezvous store (``request_resize``); the ring is re-formed with a fresh ncclUniqueId.
"""
from __future__ import annotations

import copy
import functools
import pickle
import time

import torch

from . import _native

_PLAN_KEY = "ddp_step/elastic/plan"


reset_latencies_ms = []   # wall time of every elastic reset seen by this process: interrupt -> new ring -> state synced


def _hvd():
    from . import hvd
    return hvd


# -- membership plan (stands in for horovod's host-discovery driver) ---------------------------------
def request_resize(ranks):
    """Announce that the next ring should consist of ``ranks`` (global ranks).  Any member may call it; every
    member notices at its next ``state.commit()`` (all at the same commit, agreed by an allreduce-max)."""
    h = _hvd()
    store = h._ctx.store
    if store is None:
        raise RuntimeError("request_resize needs a multi-process job")
    gen = _read_plan(store)[0] + 1
    store.set(_PLAN_KEY, pickle.dumps((gen, list(ranks))))


def _read_plan(store):
    if store is None or not store.check([_PLAN_KEY]):
        return 0, None
    return pickle.loads(bytes(store.get(_PLAN_KEY)))


class State:
    """Scalar attributes + reset callbacks (horovod.common.elastic.ObjectState semantics)."""

    def __init__(self, **kwargs):
        self._reset_callbacks = []
        self._saved_scalars = dict(kwargs)
        self._scalar_names = list(kwargs)
        for k, v in kwargs.items():
            setattr(self, k, v)
        self._known_generation = _hvd()._ctx.generation
        self._flag = None

    def register_reset_callbacks(self, callbacks):
        self._reset_callbacks.extend(callbacks)

    def on_reset(self):
        for cb in self._reset_callbacks:
            cb()

    # scalars
    def _save_scalars(self):
        self._saved_scalars = {k: copy.deepcopy(getattr(self, k)) for k in self._scalar_names}

    def _restore_scalars(self):
        for k, v in self._saved_scalars.items():
            setattr(self, k, copy.deepcopy(v))

    def _sync_scalars(self):
        h = _hvd()
        vals = h.broadcast_object({k: getattr(self, k) for k in self._scalar_names}, 0)
        for k, v in vals.items():
            setattr(self, k, v)
        self._save_scalars()

    def save(self):
        self._save_scalars()

    def restore(self):
        self._restore_scalars()

    def sync(self):
        self._sync_scalars()

    def commit(self):
        self.save()
        self.check_host_updates()

    def check_host_updates(self):
        """Raise HostsUpdatedInterrupt on EVERY member at the same commit if a resize was announced."""
        h = _hvd()
        c = h._ctx.comm
        if c is None or h._ctx.store is None:
            return
        seen = _read_plan(h._ctx.store)[0]
        dev = "cuda" if torch.cuda.is_available() else "cpu"
        if self._flag is None or self._flag.device.type != dev:
            self._flag = torch.zeros(1, dtype=torch.int64, device=dev)
        self._flag.fill_(seen)
        c.allreduce_max_(self._flag)
        agreed = int(self._flag.item())
        if agreed > self._known_generation:
            self._pending_generation = agreed
            raise h.HostsUpdatedInterrupt(skip_sync=False)


class TorchState(State):
    """ref: horovod/horovod_mnist_elastic.py:104 — ``TorchState(model, optimizer, batch=0, epoch=0)``."""

    def __init__(self, model=None, optimizer=None, **kwargs):
        super().__init__(**kwargs)
        self.model = model
        self.optimizer = optimizer
        self._host = None        # pinned host copies of the arenas
        self._host_extra = None  # deep copy of non-parameter state (buffers)
        self._event = None
        self.save()

    # -- what must be captured -------------------------------------------------------------------------
    def _arenas(self):
        fused = getattr(self.optimizer, "_fused", None)
        if fused is not None:
            for g in fused.groups:
                g.sync_step_to_device()
            return fused.arenas()
        if self.model is not None and getattr(self.model, "_arena", None) is not None:
            return [self.model.flat_params]
        return None

    def _extra_state(self):
        """Tensors of model.state_dict() that are not parameters (e.g. BatchNorm buffers)."""
        if self.model is None:
            return {}
        pnames = {n for n, _ in self.model.named_parameters()}
        return {k: v for k, v in self.model.state_dict().items() if k not in pnames}

    def save(self):
        self._save_scalars()
        arenas = self._arenas() if (self.model is not None or self.optimizer is not None) else None
        if arenas is None:
            # generic fallback for objects we do not manage flat: plain deep copies (host logic only)
            self._generic = (copy.deepcopy(self.model.state_dict()) if self.model is not None else None,
                             copy.deepcopy(self.optimizer.state_dict()) if self.optimizer is not None else None)
            return
        nat = _native.get()
        if self._host is None or len(self._host) != len(arenas) or any(
                h.numel() * h.element_size() < a.numel() * a.element_size() for h, a in zip(self._host, arenas)):
            self._host = [torch.empty(a.shape, dtype=a.dtype, pin_memory=a.is_cuda) for a in arenas]
        for a, hbuf in zip(arenas, self._host):
            nat.state_snapshot(a, hbuf)
        if torch.cuda.is_available() and arenas and arenas[0].is_cuda:
            self._event = torch.cuda.Event()
            self._event.record()
        self._host_extra = {k: v.detach().to("cpu", copy=True) for k, v in self._extra_state().items()}

    def restore(self):
        self._restore_scalars()
        arenas = self._arenas() if (self.model is not None or self.optimizer is not None) else None
        if arenas is None:
            ms, os_ = self._generic
            if ms is not None:
                self.model.load_state_dict(ms)
            if os_ is not None:
                self.optimizer.load_state_dict(os_)
            return
        if self._event is not None:
            self._event.synchronize()
        nat = _native.get()
        for a, hbuf in zip(arenas, self._host):
            nat.state_restore(a, hbuf)
        fused = getattr(self.optimizer, "_fused", None)
        if fused is not None:
            torch.cuda.current_stream().synchronize() if torch.cuda.is_available() else None
            fused.set_steps_from_device()
        if self._host_extra:
            sd = self.model.state_dict()
            for k, v in self._host_extra.items():
                sd[k].copy_(v)

    def sync(self):
        """Broadcast model + optimizer + scalars from rank 0 (what @hvd.elastic.run does before each attempt)."""
        h = _hvd()
        if self.model is not None:
            h.broadcast_parameters(self.model.state_dict(), root_rank=0)
        if self.optimizer is not None and getattr(self.optimizer, "_fused", None) is not None:
            h.broadcast_optimizer_state(self.optimizer, root_rank=0)
        self._sync_scalars()
        self.save()


def _reset(state):
    """hvd.shutdown(); hvd.init() of horovod's elastic reset: adopt the announced membership."""
    h = _hvd()
    gen, ranks = _read_plan(h._ctx.store)
    if ranks is None:
        ranks = list(range(h._ctx.job_world))
    h.reform(ranks, generation=gen)
    state._known_generation = gen
    # a process left out of the ring idles until a later plan lists it again, then rejoins
    while h._ctx.comm is None:
        time.sleep(0.05)
        g2, r2 = _read_plan(h._ctx.store)
        if g2 > gen and r2 is not None:
            gen = g2
            h.reform(r2, generation=g2)
            state._known_generation = g2


def run(func):
    """ref: horovod/horovod_mnist_elastic.py:55 — decorator that makes ``func(state, ...)`` elastic."""

    @functools.wraps(func)
    def wrapper(state, *args, **kwargs):
        h = _hvd()
        skip_sync = False
        t_interrupt = None
        while True:
            if not skip_sync:
                state.sync()
            if t_interrupt is not None:          # membership change -> ring re-formed -> state broadcast done
                if torch.cuda.is_available():
                    torch.cuda.synchronize()
                reset_latencies_ms.append((time.perf_counter() - t_interrupt) * 1e3)
                t_interrupt = None
            try:
                return func(state, *args, **kwargs)
            except h.HorovodInternalError:
                state.restore()
                skip_sync = False
            except h.HostsUpdatedInterrupt as e:
                skip_sync = e.skip_sync
            t_interrupt = time.perf_counter()
            _reset(state)
            state.on_reset()

    return wrapper


def __getattr__(name):   # hvd.elastic.HorovodInternalError etc.
    if name in ("HorovodInternalError", "HostsUpdatedInterrupt"):
        return getattr(_hvd(), name)
    raise AttributeError(name)
