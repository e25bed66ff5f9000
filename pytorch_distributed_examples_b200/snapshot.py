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
