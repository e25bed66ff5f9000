"""``TrainStep`` — the whole data-parallel training step of the reference loop as ONE replayable CUDA graph.

The reference's hot loop (ref: horovod/mnist_horovod.py:59-64; Trainer._run_batch,
ref: pytorch_elastic/mnist_ddp_elastic.py:71-79) is
    optimizer.zero_grad(); output = model(data); loss = F.nll_loss(output, target); loss.backward(); optimizer.step()
At 2.6 MFLOP/sample the GPU work of a step is tens of microseconds, so Python dispatch, autograd and launch
latency dominate unless the step is fused and replayed.  TrainStep enqueues, on one stream,
    K_F1..K_F3 (+NLL mean)  ->  K_B3,K_B2d,K_B2w,K_B1,K_R (grads born in the bucket)
    ->  ncclAllReduce(sum) of the bucket (world > 1)  ->  fused (1/N)*g + optimizer update  ->  step++
captures it once, and replays it per batch.  zero_grad() is free (the bucket is overwritten).  Dropout masks
come from Philox keyed by the device step counter, so replays draw fresh masks.
Same arithmetic as the module path (`Net` + `hvd.DistributedOptimizer`), which stays the drop-in API.

``slots`` > 1 keeps several resident input batches (one captured graph each) so a benchmark can rotate through
more input bytes than the L2 holds.
"""
from __future__ import annotations

import os

import torch

from . import _native
from .net import Net
from .optim import FusedUpdater

KERNELS_PER_STEP = 2 + 1 + 4 + 1   # small-batch chain: conv forward x2, fused head, backward x3 + reduce, fused update (also advances the
                                   # step counter); with the head as two kernels (B > 2048) it is 9 -- TrainStep.kernel_launches_per_step


class TrainStep:
    def __init__(self, model: Net, optimizer, batch_size: int, communicator=None, use_graph: bool = True,
                 slots: int = 1, collective: str = "auto"):
        model._check_arena()
        self.model = model
        self.optimizer = optimizer
        self.updater = getattr(optimizer, "_fused", None) or FusedUpdater(optimizer)
        if len(self.updater.groups) != 1 or self.updater.groups[0].param_arena.data_ptr() != model.flat_params.data_ptr():
            raise RuntimeError("TrainStep expects one param group covering exactly the Net's parameters")
        self.group = self.updater.groups[0]
        self.comm = communicator
        self.world = communicator.world if communicator is not None else 1
        # the one exchange of the step: "peer" = one-shot NVLink peer-memory allreduce fused with the update
        # (latency ~10 us, one node, <= 8 ranks), "nccl" = ncclAllReduce + fused update, "auto" = peer if possible
        if collective not in ("auto", "nccl", "peer"):
            raise ValueError(collective)
        self.use_peer = False
        if self.world > 1 and collective != "nccl":
            self.use_peer = communicator.enable_peer(_native.ARENA_ELEMS)
            if collective == "peer" and not self.use_peer:
                raise RuntimeError("peer-memory allreduce needs all ranks on one node with P2P access (world <= 8)")
        self.batch = int(batch_size)
        dev = model.flat_params.device
        nat = _native.get()
        self.xs = [torch.zeros(self.batch, 1, 28, 28, dtype=torch.float32, device=dev) for _ in range(slots)]
        self.ys = [torch.zeros(self.batch, dtype=torch.int64, device=dev) for _ in range(slots)]
        self.logp = torch.empty(self.batch, 10, dtype=torch.float32, device=dev)
        self.losses = [torch.zeros(1, dtype=torch.float32, device=dev) for _ in range(slots)]   # one per slot/graph
        self.loss = self.losses[0]
        self._loss_host = None             # pinned host copies + events for asynchronous loss read-back
        self._loss_events = [None] * slots
        self.ws = torch.empty(nat.workspace_bytes(self.batch), dtype=torch.uint8, device=dev)
        self.scratch = torch.zeros(nat.scratch_bytes(self.batch), dtype=torch.uint8, device=dev)
        self.copy_stream = torch.cuda.Stream(device=dev)
        self._staged = [None] * slots      # event: H2D copy into the slot finished
        self._consumed = [None] * slots    # event: last step that read the slot finished
        self._u8 = None                    # per-slot uint8 staging buffers (stage_u8)
        self.use_graph = use_graph
        # the fused head kernel (ddp_net_train_fwdbwd) removes a launch, a second staging of fc1.weight and the h / logp
        # round trip: worth 8 us of an 80 us step at B = 64; from B = 4096 up the two separately tuned kernels are
        # ~3 us faster (measured), so large batches keep them.  DDP_SPLIT_HEAD=1 / DDP_FUSED_HEAD=1 force either.
        self.fused_head = (self.batch <= 2048 or os.environ.get("DDP_FUSED_HEAD", "0") == "1") \
            and os.environ.get("DDP_SPLIT_HEAD", "0") != "1"
        # A/B switch DDP_WHOLE_STEP=1: ONE call for forward + backward + exchange + update (ddp_net_train_step_peer), the
        # exchange kernel reducing the backward's partial rows itself.  Measured at 2 GPUs, B = 4096: 256.6 us/step against
        # 255.7 us for backward -> reduce_partials -> peer allreduce + update, i.e. no gain (what stays exposed is the flag
        # round trip and the step-to-step skew between the ranks), so the longer-tested sequence stays the default.
        self.whole_step = os.environ.get("DDP_WHOLE_STEP", "0") == "1"
        # our kernels per step: conv1_fwd, conv2_fwd, head (1 fused / 2 separate), conv2_dgrad, conv2_wgrad, conv1_wgrad,
        # reduce_partials, update (fused_update, or the peer allreduce+update kernel) -- ncclAllReduce, when used, is not ours
        self.kernel_launches_per_step = 8 if self.fused_head else 9
        if self.use_peer and self.whole_step:
            self.kernel_launches_per_step -= 1           # the exchange kernel reduces the partial rows itself
        self.graphs = None
        self._fingerprint = None

    @property
    def x(self):
        return self.xs[0]

    @property
    def y(self):
        return self.ys[0]

    # -- one step's stream work ----------------------------------------------------------------------
    def _enqueue(self, slot: int = 0):
        nat = _native.get()
        m, g = self.model, self.group
        x, y = self.xs[slot], self.ys[slot]
        if self.use_peer and self.whole_step:
            # forward, loss, backward, NVLink exchange and update in ONE call: the partial gradient sums go straight into
            # the peer exchange, most of them while conv1's weight gradient is still being computed
            g.bucket = m._grad_arena
            g.adopted = True
            g.graph_mode = True
            g.ensure_state()
            g.steps += 1
            hyper = g.hyper()
            hyper.bump_step_dev = 1
            nat.net_train_step_peer(m._arena, x, y, m.dropout_seed, 0, self.ws, self.logp, self.losses[slot], self.scratch,
                                    m._grad_arena, self.comm.peer, g.kind, g.state1, g.state2, hyper, step=g.steps,
                                    step_dev=g.step_dev, offset_dev=g.step_dev, fused_head=self.fused_head)
            return
        if self.fused_head:      # forward + backward in one call, the head as ONE kernel (ddp_net_train_fwdbwd)
            nat.net_train_fwdbwd(m._arena, x, y, None, None, m.dropout_seed, 0, self.ws, self.logp, self.losses[slot],
                                 self.scratch, m._grad_arena, offset_dev=g.step_dev)
        else:
            nat.net_forward(m._arena, x, y, None, None, m.dropout_seed, 0, True, self.ws, self.logp,
                            out_loss=self.losses[slot], scratch=self.scratch, offset_dev=g.step_dev)
            nat.net_backward(m._arena, x, self.logp, None, y, 1.0, True, self.ws, self.scratch, m._grad_arena)
        g.bucket = m._grad_arena
        g.adopted = True
        if self.use_peer:
            g.update_peer(self.comm, use_step_dev=True)
        else:
            if self.world > 1:
                self.comm.allreduce_sum_(m._grad_arena)
            g.update(1.0 / self.world, use_step_dev=True)

    def _hyper_fingerprint(self):
        return tuple(sorted((k, str(v)) for k, v in self.group.group.items() if k != "params"))

    def _capture(self):
        g = self.group
        g.ensure_state()
        g.sync_step_to_device()
        keep = [t.clone() for t in (g.param_arena, g.step_dev, self.losses[0])]
        keep_s = [s.clone() if s is not None else None for s in (g.state1, g.state2)]
        host_steps = g.steps
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):            # warm-up outside capture (sets func attributes, loads modules)
                self._enqueue(0)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        with torch.no_grad():             # undo the warm-up steps
            g.param_arena.copy_(keep[0]); g.step_dev.copy_(keep[1]); self.losses[0].copy_(keep[2])
            for s, k in zip((g.state1, g.state2), keep_s):
                if s is not None:
                    s.copy_(k)
        graphs = []
        for slot in range(len(self.xs)):
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                self._enqueue(slot)
            graphs.append(graph)
        g.steps = host_steps
        self.graphs = graphs
        self._fingerprint = self._hyper_fingerprint()

    # -- public --------------------------------------------------------------------------------------
    def stage(self, slot: int, data: torch.Tensor, target: torch.Tensor):
        """Prefetch a (pinned) host batch into ``slot`` on the copy stream while earlier steps compute —
        what DataLoader(pin_memory=True) + non_blocking copies do in the reference-style loop."""
        cs = self.copy_stream
        if self._consumed[slot] is not None:
            cs.wait_event(self._consumed[slot])
        with torch.cuda.stream(cs):
            self.xs[slot].copy_(data.reshape(self.batch, 1, 28, 28), non_blocking=True)
            self.ys[slot].copy_(target, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(cs)
        self._staged[slot] = ev

    def stage_u8(self, slot: int, pixels_u8: torch.Tensor, target: torch.Tensor):
        """Like ``stage`` for a raw uint8 batch [B,784] (MNIST's storage format): 4x fewer PCIe bytes; the
        ToTensor+Normalize of the reference's transform (ref: horovod/mnist_horovod.py:34-37) runs on the
        device (ddp_normalize_u8) on the copy stream."""
        cs = self.copy_stream
        if self._u8 is None:
            self._u8 = [torch.empty(self.batch * 784, dtype=torch.uint8, device=self.xs[0].device) for _ in self.xs]
        if self._consumed[slot] is not None:
            cs.wait_event(self._consumed[slot])
        with torch.cuda.stream(cs):
            self._u8[slot].copy_(pixels_u8.reshape(-1), non_blocking=True)
            self.ys[slot].copy_(target, non_blocking=True)
            _native.get().normalize_u8(self._u8[slot], self.xs[slot])
            ev = torch.cuda.Event()
            ev.record(cs)
        self._staged[slot] = ev

    def run_slot(self, slot: int = 0) -> torch.Tensor:
        """One step on the batch resident (or being staged) in ``xs[slot]`` / ``ys[slot]``."""
        if self.use_graph and (self.graphs is None or self._fingerprint != self._hyper_fingerprint()):
            self._capture()
        if self._staged[slot] is not None:
            torch.cuda.current_stream().wait_event(self._staged[slot])
            self._staged[slot] = None
        if not self.use_graph:
            self._enqueue(slot)
        else:
            self.graphs[slot].replay()
            self.group.steps += 1
            self.group.graph_mode = True
        if len(self.xs) > 1:
            ev = torch.cuda.Event()
            ev.record()
            self._consumed[slot] = ev
        self.loss = self.losses[slot]
        return self.loss

    def post_loss(self, slot: int):
        """Queue an asynchronous device->host copy of the loss of the step just enqueued on ``slot`` (pinned
        memory + event), so the host can read step i's loss while step i+1 is already running."""
        if self._loss_host is None:
            self._loss_host = torch.zeros(len(self.xs), dtype=torch.float32).pin_memory()
            self._loss_stream = torch.cuda.Stream(device=self.xs[0].device)
        done = torch.cuda.Event()
        done.record()                                  # the step just enqueued on the compute stream
        ls = self._loss_stream
        ls.wait_event(done)
        with torch.cuda.stream(ls):                    # off the compute stream: the next step is not queued behind the copy
            self._loss_host[slot:slot + 1].copy_(self.losses[slot], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(ls)
        self._loss_events[slot] = ev

    def read_loss(self, slot: int) -> float:
        """The loss posted by ``post_loss(slot)``; waits only for that step, not for later ones."""
        self._loss_events[slot].synchronize()
        return float(self._loss_host[slot])

    def __call__(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        """One step on (data [B,1,28,28], target [B]); returns the device loss tensor (no host sync).
        Host tensors are accepted (pinned memory makes the copies asynchronous)."""
        if data.shape[0] != self.batch:
            raise RuntimeError(f"TrainStep was built for batch {self.batch}, got {data.shape[0]}")
        if data.data_ptr() != self.xs[0].data_ptr():
            self.xs[0].copy_(data.reshape(self.batch, 1, 28, 28), non_blocking=True)
        if target.data_ptr() != self.ys[0].data_ptr():
            self.ys[0].copy_(target, non_blocking=True)
        return self.run_slot(0)
