"""DDP communication hook: stock ``torch.nn.parallel.DistributedDataParallel`` + our NCCL path.

The reference wraps its model in DDP (ref: pytorch_elastic/mnist_ddp_elastic.py:58) and relies on DDP's
built-in bucket allreduce.  ``allreduce_hook`` is the drop-in seam: registered with
``ddp.register_comm_hook(state, allreduce_hook)`` it must leave results equal to hook-less DDP, whose
contract is torch: distributed/algorithms/ddp_comm_hooks/default_hooks.py:35-54 — the hook receives the
bucket NOT pre-divided and returns a Future of a tensor shaped like ``bucket.buffer()`` holding the mean.
Here: ncclAllReduce(sum) of the flat bucket through the C-ABI on a dedicated comm stream, then the 1/N scale
kernel, completed as a CUDA-aware Future so DDP's finalize_backward stream-syncs correctly.
``FusedUpdateHook`` goes one step further (optimizer-in-backward, same idea as torch:
distributed/algorithms/ddp_comm_hooks/optimizer_overlap_hooks.py:50-128): after the allreduce it runs the
fused (1/N)*g + optimizer update straight from the bucket, so ``optimizer.step()`` becomes a no-op.
"""

import torch
import torch.distributed as dist

from . import _native
from .optim import optimizer_kind


class HookState:
    """``symmetric_bytes`` > 0 (collective: pass the same value on every rank) additionally sets up the two-shot
    NVSwitch allreduce (csrc/sym_allreduce.cu) for buckets of at least ``symmetric_min_bytes`` and at most
    ``symmetric_bytes``: the bucket is staged through a symmetric buffer, reduced in the switch, and handed
    back averaged.  Smaller / larger buckets, and boxes without NVLink symmetric memory, use ncclAllReduce."""

    def __init__(self, communicator=None, device=None, symmetric_bytes: int = 0, symmetric_min_bytes: int = 1 << 20):
        if communicator is None:
            from . import hvd
            communicator = hvd._require()
        self.comm = communicator
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device, priority=-1)
        self.calls = 0
        self.sym_calls = 0
        self.symmetric = bool(symmetric_bytes) and communicator.enable_symmetric(int(symmetric_bytes))
        self.symmetric_min_bytes = int(symmetric_min_bytes)


def _average_bucket(state: HookState, buf: torch.Tensor):
    """buf <- mean over ranks, on the current (comm) stream."""
    comm, world = state.comm, state.comm.world
    nbytes = buf.numel() * buf.element_size()
    if state.symmetric and state.symmetric_min_bytes <= nbytes <= comm.sym["capacity"] \
            and buf.dtype in (torch.float32, torch.bfloat16):
        per16 = 16 // buf.element_size()
        n = buf.numel()
        n_pad = -(-n // per16) * per16
        v = comm.sym_view(buf.dtype, n_pad)
        v[:n].copy_(buf)
        if n_pad != n:
            v[n:].zero_()
        comm.sym_allreduce_(buf.dtype, n_pad, 0, 1.0 / world)
        buf.copy_(v[:n])
        state.sym_calls += 1
        return
    comm.allreduce_sum_(buf)
    if world > 1:
        _native.get().scale(buf, 1.0 / world)


def _run_on_comm_stream(state: HookState, buf: torch.Tensor, body):
    nat = _native.get()
    nat.bind_device(state.device.index)        # hooks run on autograd's engine thread
    cur = torch.cuda.current_stream(state.device)
    state.stream.wait_stream(cur)
    with torch.cuda.stream(state.stream):
        body()
        fut = torch.futures.Future(devices=[state.device])
        fut.set_result(buf)                    # records an event on the comm stream
    buf.record_stream(state.stream)
    state.calls += 1
    return fut


def allreduce_hook(state: HookState, bucket: dist.GradBucket) -> torch.futures.Future[torch.Tensor]:
    buf = bucket.buffer()
    return _run_on_comm_stream(state, buf, lambda: _average_bucket(state, buf))


class FusedUpdateHook:
    """``ddp.register_comm_hook(FusedUpdateHook.State(optimizer), FusedUpdateHook.hook)``.
    Parameters of each bucket are re-homed into a flat arena packed exactly like DDP packs the bucket, so
    the update kernel walks bucket and parameters with the same index.  DDP rebuilds its buckets after the
    first iteration (backward order); optimizer state is therefore tracked per PARAMETER and carried into
    the new packing when a bucket's composition changes."""

    class State(HookState):
        def __init__(self, optimizer: torch.optim.Optimizer, communicator=None, device=None, **kw):
            super().__init__(communicator, device, **kw)
            self.optimizer = optimizer
            self.kind = optimizer_kind(optimizer)
            self.buckets = {}       # bucket index -> dict(key, arena, state1, state2, shell)
            self.param_state = {}   # id(param) -> (state1 view, state2 view)
            self.iteration = 0      # completed backward passes == optimizer steps taken

    @staticmethod
    def _prepare(state, bucket):
        from .optim import _Group
        params = bucket.parameters()
        total = bucket.buffer().numel()
        group = None
        for p in params:
            for g in state.optimizer.param_groups:
                if any(p is q for q in g["params"]):
                    if group is not None and g is not group:
                        raise RuntimeError("a DDP bucket spans several optimizer param groups")
                    group = g
        shell = _Group.__new__(_Group)      # reuse hyper()/needs_state() on a bare group
        shell.group, shell.kind = group, state.kind
        n1, n2 = shell.needs_state()
        arena = torch.zeros(total, dtype=torch.float32, device=state.device)
        s1 = torch.zeros_like(arena) if n1 else None
        s2 = torch.zeros_like(arena) if n2 else None
        off = 0
        with torch.no_grad():
            for p in params:
                n = p.numel()
                arena[off:off + n].copy_(p.data.reshape(-1))
                p.data = arena[off:off + n].view(p.shape)
                old = state.param_state.get(id(p))
                if old is not None:             # carry optimizer state across DDP's bucket rebuild
                    if s1 is not None and old[0] is not None:
                        s1[off:off + n].copy_(old[0])
                    if s2 is not None and old[1] is not None:
                        s2[off:off + n].copy_(old[1])
                state.param_state[id(p)] = (s1[off:off + n] if s1 is not None else None,
                                            s2[off:off + n] if s2 is not None else None)
                off += n
        return dict(key=tuple(id(p) for p in params), arena=arena, shell=shell, state1=s1, state2=s2)

    @staticmethod
    def hook(state, bucket: dist.GradBucket) -> torch.futures.Future[torch.Tensor]:
        buf = bucket.buffer()
        idx = bucket.index()
        key = tuple(id(p) for p in bucket.parameters())
        if idx not in state.buckets or state.buckets[idx]["key"] != key:
            state.buckets[idx] = FusedUpdateHook._prepare(state, bucket)
        b = state.buckets[idx]
        world = state.comm.world
        step = state.iteration + 1
        if bucket.is_last():
            state.iteration += 1

        def body():
            hyper = b["shell"].hyper(write_back=True)   # leave the averaged gradient in the bucket for DDP
            if state.symmetric:
                _average_bucket(state, buf)
                inv = 1.0
            else:
                state.comm.allreduce_sum_(buf)
                inv = 1.0 / world
            _native.get().fused_update(state.kind, buf, b["arena"], b["state1"], b["state2"], buf.numel(),
                                       inv, hyper, step=step)

        return _run_on_comm_stream(state, buf, body)


class BufferBroadcast:
    """DDP's ``broadcast_buffers=True`` through OUR communicator.

    Stock DDP re-broadcasts every module buffer (BatchNorm ``running_mean`` / ``running_var`` /
    ``num_batches_tracked`` of the ResNet-50 of ref: rpc/model_parallel_ResNet50.py:94-101,126-132) from rank 0 before
    each forward, as coalesced c10d broadcasts (torch: nn/parallel/distributed.py ``_sync_buffers``).  Here the
    buffers are re-homed ONCE into one flat arena per dtype (same trick as the parameter arena: ``buffer.data``
    becomes a view, ``state_dict`` is unchanged) and a forward pre-hook issues ONE ncclBroadcast per arena through
    the C-ABI on the current stream -- ResNet-50: 161 buffers -> 2 broadcasts.  Use with
    ``DDP(module, broadcast_buffers=False)``::

        bb = BufferBroadcast(module)            # before or after the DDP wrap
        ddp = DDP(module, device_ids=[dev], broadcast_buffers=False)
    """

    def __init__(self, module: torch.nn.Module, communicator=None, root_rank: int = 0):
        if communicator is None:
            from . import hvd
            communicator = hvd._require()
        self.comm = communicator
        self.root = int(root_rank)
        by_dtype = {}
        for b in module.buffers():
            if b is not None and b.numel():
                by_dtype.setdefault((b.dtype, b.device), []).append(b)
        self.arenas = []
        with torch.no_grad():
            for (dtype, device), bufs in by_dtype.items():
                esz = bufs[0].element_size()
                per16 = max(1, 16 // esz)
                offs, total = [], 0
                for b in bufs:                        # every buffer starts 16-byte aligned inside the arena
                    offs.append(total)
                    total += -(-b.numel() // per16) * per16
                arena = torch.zeros(total, dtype=dtype, device=device)
                for b, o in zip(bufs, offs):
                    arena[o:o + b.numel()].copy_(b.reshape(-1))
                    b.data = arena[o:o + b.numel()].view(b.shape)
                self.arenas.append(arena)
        self.calls = 0
        self.handle = module.register_forward_pre_hook(self._pre_forward)

    def _pre_forward(self, module, inputs):
        self.sync()

    def sync(self):
        """Every rank's buffers <- the root rank's (one broadcast per dtype arena, on the current stream)."""
        if self.comm.world > 1:
            for a in self.arenas:
                self.comm.broadcast_(a, self.root)
        self.calls += 1

    def remove(self):
        self.handle.remove()
