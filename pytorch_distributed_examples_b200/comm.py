"""Process-group plumbing for the native NCCL communicator.

``torch.distributed`` is used only for rendezvous (its TCP store carries the 128-byte ncclUniqueId, small
pickled objects and the elastic membership plan); every tensor collective on the hot path goes through the
C-ABI (ddp_allreduce_sum / ddp_broadcast in csrc/comm.cu) on the caller's CUDA stream, so it can be captured
in a CUDA graph together with the kernels around it.
Replaces ref: pytorch_elastic/mnist_ddp_elastic.py:22-27 (``init_process_group("gloo")``) and ``hvd.init()``
(ref: horovod/mnist_horovod.py:28).
"""
from __future__ import annotations

import collections
import os
import pickle
import threading
import time

import torch
import torch.distributed as dist

from . import _native


def env_world():
    """(world, rank, local_rank) from torchrun / horovodrun / mpirun environments; (1,0,0) if none."""
    for w, r, l in (("WORLD_SIZE", "RANK", "LOCAL_RANK"), ("HOROVOD_SIZE", "HOROVOD_RANK", "HOROVOD_LOCAL_RANK"),
                    ("OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_LOCAL_RANK")):
        if w in os.environ and r in os.environ:
            return int(os.environ[w]), int(os.environ[r]), int(os.environ.get(l, os.environ[r]))
    return 1, 0, 0


def default_store(world: int):
    """The rendezvous key-value store.  Initialises torch.distributed (env://) if the script has not."""
    if world <= 1:
        return None
    if not dist.is_initialized():
        backend = "nccl" if torch.cuda.is_available() else "gloo"
        dist.init_process_group(backend=backend)
    return dist.distributed_c10d._get_default_store()


class Communicator:
    """One NCCL communicator over ``ranks`` (global ranks of the job), created through the store."""

    def __init__(self, store, global_rank: int, ranks, generation="0.0"):
        self.store = store
        self.global_rank = global_rank
        self.ranks = list(ranks)
        self.world = len(self.ranks)
        self.rank = self.ranks.index(global_rank)
        self.generation = generation
        self.handle = None
        self.peer = None          # NVLink peer-memory context (enable_peer)
        self.peer_max_elems = 0
        self.sym = None           # symmetric buffer for the two-shot large-bucket allreduce (enable_symmetric)
        self.sym_error = None
        self._obj_seq = 0
        self.broken = None        # reason string once the ring is known to be dead (a peer is lost); see wait()
        self.timeout_s = float(os.environ.get("DDP_COLLECTIVE_TIMEOUT_S", "120"))
        self._pending = collections.deque()   # (event, enqueue time) of collectives enqueued from Python
        self._watchdog = None
        nat = _native.get()
        if self.world > 1:
            key = f"ddp_step/uid/gen{generation}"
            if self.rank == 0:
                uid = nat.comm_unique_id()
                store.set(key, uid)
            else:
                uid = bytes(store.get(key))
                store.add(key + "/acks", 1)
            self.handle = nat.comm_init(uid, self.world, self.rank)
            if self.rank == 0:
                # Once every member has read the id, delete it: a later ring that reuses this tag
                # (hvd.shutdown(); hvd.init()) can then never pick up a stale id -- its readers block until the new one is set.
                try:
                    deadline = time.monotonic() + 60
                    while store.add(key + "/acks", 0) < self.world - 1 and time.monotonic() < deadline:
                        time.sleep(0.002)
                    store.delete_key(key)
                    store.delete_key(key + "/acks")
                except Exception:
                    pass

    # -- failure detection -----------------------------------------------------------------------------
    # A dead peer does not make an intra-node NCCL collective fail: its kernel spins forever.  Every collective
    # enqueued through this class therefore leaves a CUDA event behind; wait() polls the newest one against a deadline
    # (and ncclCommGetAsyncError for transport errors), and a watchdog thread does the same for loops whose only host
    # synchronisation is torch's own (loss.item()).  Either way a late collective ends in ncclCommAbort -- which makes
    # the stuck kernel exit -- and CollectiveError (= hvd.HorovodInternalError) at the next call into this class.
    def _note_collective(self):
        if self.world <= 1 or not torch.cuda.is_available() or torch.cuda.is_current_stream_capturing():
            return
        ev = torch.cuda.Event()
        ev.record()
        self._pending.append((ev, time.monotonic()))
        while len(self._pending) > 64:
            self._pending.popleft()
        if self._watchdog is None:
            self._watchdog = threading.Thread(target=self._watch, name="ddp-comm-watchdog", daemon=True)
            self._watchdog.start()

    def _watch(self):
        while self.handle is not None and self.broken is None:
            time.sleep(min(1.0, max(0.05, self.timeout_s / 10)))
            try:
                while self._pending and self._pending[0][0].query():
                    self._pending.popleft()
                if self._pending and time.monotonic() - self._pending[0][1] > self.timeout_s:
                    self.abort(f"a collective did not finish within {self.timeout_s:.0f} s (peer lost?)")
            except Exception:
                return

    def abort(self, reason: str):
        """Tear the ring down without waiting for it (ncclCommAbort): in-flight collective kernels exit."""
        if self.broken is None:
            self.broken = reason
        h, self.handle = self.handle, None
        if h is not None:
            try:
                _native.get().comm_abort(h)
            except Exception:
                pass

    def raise_if_broken(self):
        if self.broken is not None:
            raise _native.CollectiveError(self.broken)

    def wait(self, timeout_s: float | None = None):
        """Block until every collective enqueued so far has finished; CollectiveError if the ring is broken, reports an
        asynchronous NCCL error, or the deadline passes (the ring is aborted first, so the stream drains)."""
        self.raise_if_broken()
        if self.world <= 1 or not torch.cuda.is_available():
            return
        timeout_s = self.timeout_s if timeout_s is None else timeout_s
        ev = torch.cuda.Event()
        ev.record()
        t0 = time.monotonic()
        nat = _native.get()
        while not ev.query():
            if self.broken is not None:
                break
            if self.handle is not None:
                err = nat.comm_async_error(self.handle)
                if err not in (0, 7):
                    self.abort(f"asynchronous NCCL error {err}")
                    break
            if time.monotonic() - t0 > timeout_s:
                self.abort(f"a collective did not finish within {timeout_s:.0f} s (peer lost?)")
                break
            time.sleep(0.0005)
        self.raise_if_broken()

    # -- NVLink peer-memory path for latency-bound buckets -------------------------------------------
    def enable_peer(self, max_elems: int) -> bool:
        """Set up the one-shot peer-memory allreduce (csrc/peer.cu) for buckets of up to ``max_elems`` floats.
        Needs every rank on ONE node with NVLink/P2P IPC (world <= 8); returns False (and NCCL stays the path)
        when the topology does not allow it.  Collective: every member must call it."""
        if self.world <= 1 or self.world > 8 or not torch.cuda.is_available():
            return False
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
        if local_world < self.world:
            return False              # ranks spread over several nodes: no shared NVLink domain
        if self.peer is not None and self.peer_max_elems >= max_elems:
            return True
        nat = _native.get()
        max_elems = (int(max_elems) + 3) // 4 * 4
        if self.peer is not None:                       # growing: release the old context (device memory + IPC mappings)
            torch.cuda.synchronize()
            nat.peer_destroy(self.peer)
            self.peer, self.peer_max_elems = None, 0
        self._peer_seq = getattr(self, "_peer_seq", 0) + 1
        peer = nat.peer_create(self.world, self.rank, max_elems)
        base = f"ddp_step/peer/gen{self.generation}/{self._peer_seq}/{max_elems}"
        self.store.set(f"{base}/{self.rank}", nat.peer_local_handle(peer))
        handles = b"".join(bytes(self.store.get(f"{base}/{r}")) for r in range(self.world))
        nat.peer_connect(peer, handles)
        # nobody may push before everyone has mapped everyone: rendezvous once through NCCL
        flag = torch.zeros(1, device="cuda")
        self.allreduce_sum_(flag)
        self.wait()
        try:                                            # every member has read every handle: no stale keys for a later ring
            self.store.delete_key(f"{base}/{self.rank}")
        except Exception:
            pass
        self.peer, self.peer_max_elems = peer, max_elems
        return True

    def enable_symmetric(self, nbytes: int, multicast: bool = True) -> bool:
        """Set up the two-shot allreduce for LARGE buckets (csrc/sym_allreduce.cu): a symmetric buffer of
        ``nbytes`` (+ flag block) on every rank, each replica mapped on every peer and, when the NVSwitch
        supports it and ``multicast`` is True, bound to one multicast object so the reduction runs in the switch.
        The memory comes from torch's symmetric-memory allocator (device-memory plumbing) and therefore needs an
        initialised torch.distributed NCCL group containing exactly this communicator's ranks, one node,
        world <= 8.  Returns False (NCCL stays the path) when any of that is missing.  Collective."""
        import torch.distributed as dist
        if self.world <= 1 or self.world > 8 or not torch.cuda.is_available():
            return False
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() != self.world \
                or dist.get_rank() != self.rank:
            return False
        if self.sym is not None and self.sym["capacity"] >= nbytes:
            return True
        nat = _native.get()
        try:
            import torch.distributed._symmetric_memory as symm_mem
            cap = (int(nbytes) + 127) // 128 * 128
            total = cap + nat.sym_flag_bytes()
            dev = torch.device("cuda", torch.cuda.current_device())
            buf = symm_mem.empty(total, dtype=torch.uint8, device=dev)
            buf.zero_()
            hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
            torch.cuda.synchronize()
            hdl.barrier()                  # every replica is zeroed before anyone's first flag store can land
            torch.cuda.synchronize()
        except Exception as exc:           # no NVLink P2P / unsupported platform: stay on NCCL
            self.sym_error = repr(exc)
            return False
        # per-direction NVLink traffic: (1 + 1/W) S through the switch vs 2 (W-1)/W S with peer loads/stores,
        # so the multicast path only pays from 4 ranks up (measured: tools/bucket_bench.py at W=2 and W=8)
        mc = int(hdl.multicast_ptr) if (multicast and self.world >= 4 and getattr(hdl, "multicast_ptr", 0)) else 0
        self.sym = dict(buf=buf, hdl=hdl, capacity=cap, flag_offset=cap, multicast=mc,
                        multicast_available=int(getattr(hdl, "multicast_ptr", 0) or 0), replicas=[int(p) for p in hdl.buffer_ptrs])
        return True

    def sym_view(self, dtype: torch.dtype, n: int, byte_offset: int = 0) -> torch.Tensor:
        """The local replica's elements [byte_offset, byte_offset + n) as a tensor: fill it with this rank's
        gradients, call ``sym_allreduce_``, read the average back from the same tensor."""
        esz = torch.empty(0, dtype=dtype).element_size()
        if byte_offset % 16 or byte_offset + n * esz > self.sym["capacity"]:
            raise RuntimeError("sym_view: range outside the symmetric buffer (or offset not 16-byte aligned)")
        return self.sym["buf"][byte_offset:byte_offset + n * esz].view(dtype)

    def sym_allreduce_(self, dtype: torch.dtype, n: int, byte_offset: int = 0, scale: float = 1.0, max_ctas: int = 0, unroll: int = 0):
        """In place on the current stream: elements <- scale * sum over ranks (all replicas end bit-identical)."""
        s = self.sym
        _native.get().sym_allreduce(s["replicas"], s["multicast"], s["flag_offset"], self.world, self.rank,
                                    byte_offset, n, dtype, scale, max_ctas, unroll)

    def peer_allreduce_(self, t: torch.Tensor):
        _native.get().peer_allreduce(self.peer, t)
        return t

    # -- tensor collectives (in place, on the current stream) ---------------------------------------
    def allreduce_sum_(self, t: torch.Tensor):
        if self.world > 1:
            self.raise_if_broken()
            _native.get().allreduce_sum(self.handle, t)
            self._note_collective()
        return t

    def allreduce_max_(self, t: torch.Tensor):
        if self.world > 1:
            self.raise_if_broken()
            _native.get().allreduce_max(self.handle, t)
            self._note_collective()
        return t

    def broadcast_(self, t: torch.Tensor, root: int = 0):
        if self.world > 1:
            self.raise_if_broken()
            _native.get().broadcast(self.handle, t, root)
            self._note_collective()
        return t

    # -- small host objects through the store --------------------------------------------------------
    def broadcast_object(self, obj, root: int = 0):
        if self.world <= 1:
            return obj
        self._obj_seq += 1
        key = f"ddp_step/obj/gen{self.generation}/{self._obj_seq}"
        if self.rank == root:
            self.store.set(key, pickle.dumps(obj))
            return obj
        return pickle.loads(bytes(self.store.get(key)))

    def destroy(self):
        """Orderly teardown of a healthy ring; a broken one (see abort) only releases its local resources."""
        if self.peer is not None:
            _native.get().peer_destroy(self.peer)
            self.peer = None
        self.sym = None
        h, self.handle = self.handle, None
        self._pending.clear()
        if h is not None:
            if self.broken is None:
                _native.get().comm_destroy(h)
            else:
                try:
                    _native.get().comm_abort(h)
                except Exception:
                    pass
