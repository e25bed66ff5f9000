"""Evaluation path: the reference's ``Trainer.test`` / ``test()``
(ref: pytorch_elastic/mnist_ddp_elastic.py:117-130; horovod/horovod_mnist_elastic.py:85-102).

The reference loops over the test batches in eval mode under ``torch.no_grad()``, takes ``torch.max(outputs, 1)``,
counts ``(predicted == labels).sum().item()`` — one host sync per batch — and prints a PER-RANK accuracy over that
rank's shard.  Here the forward runs in eval mode through the C-ABI with the argmax and the correct-count fused into
the head kernel (``ddp_net_forward(..., out_correct)``: one device counter, no host sync inside the loop), and the
(correct, total) pair is summed across ranks with one int64 allreduce, so every rank reports the accuracy of the
WHOLE test set.  ``Model`` (the MLP) has no fused counter: its logits' argmax is compared on the device.
"""
import torch

from .net import Net


@torch.no_grad()
def evaluate(model, batches, communicator=None):
    """``batches`` yields (images, labels) device tensors (e.g. a DeviceLoader over the rank's DistributedSampler
    shard).  Returns (correct, total, accuracy) over all ranks of ``communicator`` (this rank only when None)."""
    was_training = model.training
    model.eval()
    dev = next(model.parameters()).device
    counts = torch.zeros(2, dtype=torch.int64, device=dev)           # [correct, total]
    try:
        for images, labels in batches:
            if isinstance(model, Net):
                model.count_correct(images, labels, counts[0:1])     # accumulates into the device counter
            else:
                predicted = model(images).argmax(1)                   # torch.max(outputs, 1)[1], ref :127
                counts[0] += (predicted == labels).sum()
            counts[1] += labels.shape[0]
        if communicator is not None and communicator.world > 1:
            communicator.allreduce_sum_(counts)
        correct, total = (int(v) for v in counts.tolist())           # the one host sync
    finally:
        model.train(was_training)
    return correct, total, (correct / total if total else float("nan"))
