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
_ALIVE_KEY = "ddp_step/elastic/alive"


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

    # scalars.  Commits are double-buffered: a commit is STAGED into the spare copy, then the ring's health is checked
    # (with a deadline), and only then does the spare become "the last commit".  A commit taken while a collective of a
    # lost peer was still stuck on the GPU would snapshot garbage; this way restore() never sees it.
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

    def _stage(self):
        self._staged_scalars = {k: copy.deepcopy(getattr(self, k)) for k in self._scalar_names}

    def _promote(self):
        self._saved_scalars = self._staged_scalars

    def save(self):
        self._stage()
        self._promote()

    def restore(self):
        self._restore_scalars()

    def sync(self):
        self._sync_scalars()

    def commit(self):
        """ref: horovod/horovod_mnist_elastic.py:72.  Stage a copy of the state, check the ring (a lost peer raises
        HorovodInternalError here and the staged copy is dropped), make the copy the last commit, then act on an
        announced membership change."""
        h = _hvd()
        self._stage()
        try:
            self.check_host_updates()
        except h.HostsUpdatedInterrupt:
            self._promote()
            raise
        self._promote()

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
        c.wait()                 # with a deadline: a lost peer surfaces here as HorovodInternalError, not as a hang
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
        self._host = None        # two sets of pinned host copies of the arenas: last commit + the one being staged
        self._host_extras = None # deep copies of non-parameter state (buffers), per set
        self._events = None      # CUDA event fencing each set's asynchronous copies
        self._good = 0           # which set is the last confirmed commit
        self._staged = False
        self._staged_generic = None
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

    def _stage(self):
        State._stage(self)
        arenas = self._arenas() if (self.model is not None or self.optimizer is not None) else None
        if arenas is None:
            # generic fallback for objects we do not manage flat: plain deep copies (host logic only)
            self._staged_generic = (copy.deepcopy(self.model.state_dict()) if self.model is not None else None,
                                    copy.deepcopy(self.optimizer.state_dict()) if self.optimizer is not None else None)
            return
        nat = _native.get()
        spare = 1 - self._good
        if self._host is None:
            self._host = [None, None]
            self._events = [None, None]
            self._host_extras = [None, None]
        hb = self._host[spare]
        if hb is None or len(hb) != len(arenas) or any(
                h.numel() * h.element_size() < a.numel() * a.element_size() for h, a in zip(hb, arenas)):
            hb = self._host[spare] = [torch.empty(a.shape, dtype=a.dtype, pin_memory=a.is_cuda) for a in arenas]
        if self._events[spare] is not None:
            self._events[spare].synchronize()          # an older copy into this buffer has long finished; be sure
        for a, hbuf in zip(arenas, hb):
            nat.state_snapshot(a, hbuf)                # asynchronous device -> pinned host copy on the current stream
        if torch.cuda.is_available() and arenas and arenas[0].is_cuda:
            ev = torch.cuda.Event()
            ev.record()
            self._events[spare] = ev
        self._host_extras[spare] = {k: v.detach().to("cpu", copy=True) for k, v in self._extra_state().items()}
        self._staged = True

    def _promote(self):
        State._promote(self)
        if getattr(self, "_staged_generic", None) is not None:
            self._generic = self._staged_generic
            self._staged_generic = None
        if getattr(self, "_staged", False):
            self._good = 1 - self._good
            self._staged = False

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
        self._staged = False                           # a staged-but-unconfirmed commit is dropped
        if self._events[self._good] is not None:
            self._events[self._good].synchronize()
        nat = _native.get()
        for a, hbuf in zip(arenas, self._host[self._good]):
            nat.state_restore(a, hbuf)
        fused = getattr(self.optimizer, "_fused", None)
        if fused is not None:
            torch.cuda.current_stream().synchronize() if torch.cuda.is_available() else None
            fused.set_steps_from_device()
        extra = self._host_extras[self._good]
        if extra:
            sd = self.model.state_dict()
            for k, v in extra.items():
                sd[k].copy_(v)

    def sync(self):
        """Broadcast model + optimizer + scalars from rank 0 (what @hvd.elastic.run does before each attempt)."""
        h = _hvd()
        if self.model is not None:
            h.broadcast_parameters(self.model.state_dict(), root_rank=0)
        if self.optimizer is not None and getattr(self.optimizer, "_fused", None) is not None:
            h.broadcast_optimizer_state(self.optimizer, root_rank=0)
        self._sync_scalars()
        c = h._ctx.comm
        if c is not None:
            c.wait()             # the broadcasts have landed (deadline: a peer lost during the sync raises here)
        self.save()


def _recover_membership(failed_comm, state):
    """After a failed collective: the survivors of the broken ring agree on who is left and publish it as the next
    membership plan (what horovod's driver does with its host-discovery script, ref: horovod_mnist_elastic.py:108).
    Every survivor registers under the broken ring's tag; the first to arrive waits a grace period for the others,
    then writes the plan; everybody adopts it in _reset()."""
    import os
    h = _hvd()
    store = h._ctx.store
    if store is None or failed_comm is None:
        return
    grace = float(os.environ.get("DDP_ELASTIC_GRACE_S", "5"))
    tag = f"{_ALIVE_KEY}/{failed_comm.generation}"
    store.set(f"{tag}/{h._ctx.global_rank}", b"1")
    first = store.add(f"{tag}/arrivals", 1) == 1
    known = state._known_generation
    if first:
        time.sleep(grace)
        alive = [r for r in failed_comm.ranks if store.check([f"{tag}/{r}"])]
        gen = max(_read_plan(store)[0], known) + 1
        store.set(_PLAN_KEY, pickle.dumps((gen, alive)))
    else:
        deadline = time.monotonic() + 10 * grace + 30
        while _read_plan(store)[0] <= known:
            if time.monotonic() > deadline:
                raise RuntimeError("elastic recovery: no membership plan was published after a failed collective")
            time.sleep(0.05)


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
                t_interrupt = time.perf_counter()
                # a peer is lost: make sure nothing of the broken ring is still running, roll back to the last
                # commit, and let the survivors agree on the next membership
                failed = h._ctx.comm
                if failed is not None:
                    failed.abort(failed.broken or "collective failed")
                if torch.cuda.is_available():
                    try:
                        torch.cuda.synchronize()
                    except RuntimeError:
                        pass
                state.restore()
                _recover_membership(failed, state)
                skip_sync = False
            except h.HostsUpdatedInterrupt as e:
                t_interrupt = time.perf_counter()
                skip_sync = e.skip_sync
            _reset(state)
            state.on_reset()

    return wrapper


def __getattr__(name):   # hvd.elastic.HorovodInternalError etc.
    if name in ("HorovodInternalError", "HostsUpdatedInterrupt"):
        return getattr(_hvd(), name)
    raise AttributeError(name)
