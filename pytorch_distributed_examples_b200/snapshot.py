"""snapshot.pt interchange with the reference (ref: pytorch_elastic/mnist_ddp_elastic.py:61-68,95-104):
``{"MODEL_STATE": model.state_dict(), "EPOCHS_RUN": epoch}`` written with ``torch.save``.  Tensors are stored
on the CPU so a snapshot written here loads in the reference's CPU-only ``torch.load(path)`` and vice versa."""
from __future__ import annotations

import torch


def save_snapshot(model: torch.nn.Module, epoch: int, path: str = "snapshot.pt", optimizer=None):
    module = model.module if hasattr(model, "module") else model     # DDP-wrapped or bare (ref :99)
    snap = {"MODEL_STATE": {k: v.detach().to("cpu") for k, v in module.state_dict().items()},
            "EPOCHS_RUN": epoch}
    if optimizer is not None:   # extension: the reference drops optimizer state (:99-102)
        snap["OPTIMIZER_STATE"] = optimizer.state_dict()
    torch.save(snap, path)


def load_snapshot(model: torch.nn.Module, path: str = "snapshot.pt", optimizer=None) -> int:
    """Returns EPOCHS_RUN (ref :66-67).  Loads before the DDP wrap, like the reference (:54-58)."""
    snap = torch.load(path, map_location="cpu")
    module = model.module if hasattr(model, "module") else model
    module.load_state_dict(snap["MODEL_STATE"])
    if optimizer is not None and "OPTIMIZER_STATE" in snap:
        optimizer.load_state_dict(snap["OPTIMIZER_STATE"])
    return snap["EPOCHS_RUN"]


class AsyncSnapshotWriter:
    """``Trainer._save_snapshot`` (ref: pytorch_elastic/mnist_ddp_elastic.py:95-104) without stalling the training
    loop: device -> pinned host on a side stream, then ``torch.save`` + atomic rename on a background thread.

    ``save(model, epoch)`` costs the training stream one event record: the side stream waits for the step that
    just finished, copies every ``state_dict`` tensor into one of TWO pinned host buffer sets (so a save can be
    staged while the previous one is still being written to disk), and a worker thread waits for that copy's
    event, builds ``{"MODEL_STATE": ..., "EPOCHS_RUN": epoch}`` from CPU clones and writes ``path + ".tmp"``
    followed by ``os.replace`` -- a reader (the reference's ``_load_snapshot``, or a relaunched job) never sees
    a torn file.  ``flush()`` waits for everything queued; errors of the worker are re-raised there and at the
    next ``save``.  CPU modules are written synchronously through ``save_snapshot``."""

    def __init__(self, path: str = "snapshot.pt"):
        import queue
        import threading
        self.path = path
        self._bufs = [None, None]            # pinned host copies, one dict per set
        self._busy = [None, None]            # threading.Event: the set has been written to disk
        self._turn = 0
        self._stream = None
        self._q = queue.Queue()
        self._err = None
        self._worker = threading.Thread(target=self._run, name="snapshot-writer", daemon=True)
        self._worker.start()

    def _run(self):
        import os
        while True:
            item = self._q.get()
            if item is None:
                return
            host, event, epoch, extra, done = item
            try:
                event.synchronize()
                snap = {"MODEL_STATE": {k: v.clone() for k, v in host.items()}, "EPOCHS_RUN": epoch}
                if extra is not None:
                    snap["OPTIMIZER_STATE"] = extra
                tmp = self.path + ".tmp"
                torch.save(snap, tmp)
                os.replace(tmp, self.path)
            except Exception as exc:          # surfaced by flush() / the next save()
                self._err = exc
            finally:
                done.set()

    def _raise_pending(self):
        if self._err is not None:
            err, self._err = self._err, None
            raise RuntimeError("asynchronous snapshot failed") from err

    def save(self, model: torch.nn.Module, epoch: int, optimizer=None):
        import threading
        self._raise_pending()
        module = model.module if hasattr(model, "module") else model
        sd = module.state_dict()
        if not any(v.is_cuda for v in sd.values()):
            save_snapshot(model, epoch, self.path, optimizer)
            return
        dev = next(v.device for v in sd.values() if v.is_cuda)
        if self._stream is None:
            self._stream = torch.cuda.Stream(device=dev)
        i = self._turn
        self._turn ^= 1
        if self._busy[i] is not None:
            self._busy[i].wait()             # this buffer set's previous snapshot is on disk
        host = self._bufs[i]
        if host is None or host.keys() != sd.keys() or any(host[k].shape != v.shape or host[k].dtype != v.dtype
                                                           for k, v in sd.items()):
            host = self._bufs[i] = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True) for k, v in sd.items()}
        extra = None
        if optimizer is not None:            # optimizer.state_dict() may sync: taken here, on the caller's thread
            extra = {k: (v if not isinstance(v, dict) else v) for k, v in optimizer.state_dict().items()}
        self._stream.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(self._stream):
            for k, v in sd.items():
                host[k].copy_(v.detach(), non_blocking=True)
                v.record_stream(self._stream)
            ev = torch.cuda.Event()
            ev.record(self._stream)
        done = threading.Event()
        self._busy[i] = done
        self._q.put((host, ev, epoch, extra, done))

    def flush(self):
        for d in self._busy:
            if d is not None:
                d.wait()
        self._raise_pending()

    def close(self):
        self.flush()
        self._q.put(None)
