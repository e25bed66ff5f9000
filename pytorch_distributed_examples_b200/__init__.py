"""pytorch_distributed_examples_b200 — the data-parallel training step of
ArnauGabrielAtienza/pytorch_distributed_examples (mnist_ddp_elastic.py, mnist_horovod.py,
horovod_mnist_elastic.py), rebuilt B200-native: hand-written sm_100a kernels behind a C-ABI
(include/ddp_step.h, csrc/), NCCL over NVLink for the one collective, CUDA graphs for the step.

Drop-in surface:
    Net                      the MNIST CNN nn.Module                      (net.py)
    Model                    the DDP script's MLP, tensor-core GEMMs      (mlp.py)
    DistributedSampler       shard indices, bit-exact with torch          (sampler.py)
    hvd                      horovod.torch-compatible module              (hvd.py, elastic.py)
    allreduce_hook / HookState   DDP register_comm_hook seam              (ddp_hook.py)
    save_snapshot / load_snapshot / AsyncSnapshotWriter  snapshot.pt interchange   (snapshot.py)
    evaluate                 Trainer.test / test(): fused count, all ranks  (evaluate.py)
plus TrainStep (the graph-captured fast path), FusedOptimizer, SyntheticMNIST / DeviceLoader.
"""
from . import _native  # noqa: F401
from .net import Net
from .mlp import Model
from .sampler import DistributedSampler
from .optim import FusedOptimizer, FusedUpdater
from .data import SyntheticMNIST, DeviceLoader
from .step import TrainStep
from .snapshot import save_snapshot, load_snapshot, AsyncSnapshotWriter
from .evaluate import evaluate
from .ddp_hook import HookState, allreduce_hook, FusedUpdateHook, BufferBroadcast
from . import hvd
from . import elastic

__all__ = ["Net", "Model", "DistributedSampler", "FusedOptimizer", "FusedUpdater", "SyntheticMNIST", "DeviceLoader",
           "TrainStep", "save_snapshot", "load_snapshot", "AsyncSnapshotWriter", "evaluate", "HookState", "allreduce_hook", "FusedUpdateHook", "BufferBroadcast",
           "hvd", "elastic"]
