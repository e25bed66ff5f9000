"""``Model`` — drop-in for the DDP script's own network (ref: pytorch_elastic/mnist_ddp_elastic.py:133-159; the
script builds ``Model(hidden_layers=5, features=1024)`` at :172 and trains it with Adam + nn.CrossEntropyLoss).

Same constructor, same attribute names (input_layer, hidden_layers, final_layer, relu) and therefore the same
``state_dict()`` keys and default initialisation.  ``model(x)`` returns the [B,10] logits wired into autograd,
so the reference loop body (ref :70-75)
    output = model(source); loss = criterion(output, targets); loss.backward(); optimizer.step()
runs unchanged, under stock DDP as well (the gradients appear in ``p.grad`` like any module's).
On a CUDA device the parameters are views of one flat fp32 arena (ddp_mlp_param_layout) and the backward
kernels write the gradients into a flat bucket of the same layout.  The wide layers are 3xTF32 tensor-core GEMMs
(csrc/mlp_gemm.cu), fp32-accurate.  The kernels want a batch that is a multiple of 4 (the weight-gradient GEMM's K is
the batch and TMA rows are 16-byte aligned): ``model(x)`` pads a ragged last batch with zero rows, which is exact
(zero rows add nothing to any gradient); ``loss_and_grads`` — the fused path, cross-entropy and its gradient
computed by the head kernels — insists on a multiple of 4.  Every autograd forward owns its activation workspace (a small
per-batch-size pool), so several forwards may be outstanding before their backwards.  No CPU implementation: CPU tensors raise.
"""
import torch
import torch.nn as nn

from . import _native


def _pad_rows(t: torch.Tensor, rows: int) -> torch.Tensor:
    """t with zero rows appended up to ``rows``.  A zero input row / zero dlogits row contributes nothing to any weight
    or bias gradient, so padding a ragged last batch (ref DataLoader, drop_last=False) to the GEMM's granule is exact."""
    if t.shape[0] == rows:
        return t
    out = torch.zeros(rows, t.shape[1], dtype=t.dtype, device=t.device)
    out[:t.shape[0]].copy_(t)
    return out


class _MLPFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, net, *params):
        nat = _native.get()
        batch = x.shape[0]
        padded = -(-batch // 4) * 4
        logits = torch.empty(padded, net._shape.out_features, dtype=torch.float32, device=x.device)
        ws = net._take_workspace(padded)          # this forward's own activations: several forwards may be outstanding
        nat.mlp_forward(net._shape, net._arena, _pad_rows(x, padded), None, ws, out_logits=logits)
        ctx.net, ctx.batch, ctx.padded, ctx.ws = net, batch, padded, ws
        net._bucket_lent = False                  # a new forward: the flat gradient bucket may be handed out again
        return logits[:batch]

    @staticmethod
    def backward(ctx, dlogits):
        nat = _native.get()
        net = ctx.net
        if ctx.ws is None:
            raise RuntimeError("Model: backward through the same forward twice (retain_graph) is not supported")
        # the flat bucket becomes p.grad of the FIRST backward node after a zero_grad(); any further node of the same
        # backward pass (two forwards feeding one loss) or an accumulation step gets a temporary and autograd adds them
        fresh = not net._bucket_lent and all(p.grad is None for p in net._param_list)
        if fresh:
            net._bucket_lent = True
        bucket = net._grad_arena if fresh else torch.empty_like(net._grad_arena)
        nat.mlp_backward(net._shape, net._arena, _pad_rows(dlogits.contiguous(), ctx.padded), None, 1.0, ctx.padded,
                         ctx.ws, bucket)
        net._give_workspace(ctx.padded, ctx.ws)   # stream-ordered: the next forward's kernels run after this backward's
        ctx.ws = None
        grads = [bucket[o:o + n].view(p.shape) for p, (o, n) in zip(net._param_list, net._layout)]
        return (None, None, *grads)


class Model(nn.Module):
    def __init__(self, hidden_layers=1, features=128):
        super().__init__()
        self.input_layer = nn.Linear(784, features)
        self.hidden_layers = nn.ModuleList([nn.Linear(features, features) for _ in range(hidden_layers)])
        self.final_layer = nn.Linear(features, out_features=10)
        self.relu = nn.ReLU()
        self._shape = _native.MLPShape(784, features, hidden_layers, 10)
        self._arena = self._grad_arena = self._workspace = None
        self._workspace_batch = 0
        self._ws_pool = {}                 # padded batch -> free activation workspaces (autograd forwards take one each)
        self._bucket_lent = False
        self._layout = None

    @property
    def _param_list(self):
        out = [self.input_layer.weight, self.input_layer.bias]
        for layer in self.hidden_layers:
            out += [layer.weight, layer.bias]
        return out + [self.final_layer.weight, self.final_layer.bias]

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        self._flatten()
        return out

    def _flatten(self):
        dev = self.input_layer.weight.device
        if dev.type != "cuda":
            self._arena = self._grad_arena = self._workspace = None
            return
        nat = _native.get()
        self._layout, total = nat.mlp_param_layout(self._shape)
        arena = torch.zeros(total, dtype=torch.float32, device=dev)
        with torch.no_grad():
            for p, (o, n) in zip(self._param_list, self._layout):
                arena[o:o + n].copy_(p.detach().reshape(-1).to(torch.float32))
                p.data = arena[o:o + n].view(p.shape)
                p.grad = None
        self._arena = arena
        self._grad_arena = torch.zeros_like(arena)
        self._workspace, self._workspace_batch = None, 0
        self._ws_pool = {}

    def _take_workspace(self, batch: int):
        """A workspace owned by ONE autograd forward until its backward returns it (a forward whose graph is dropped
        without a backward simply lets its workspace be garbage-collected)."""
        free = self._ws_pool.setdefault(batch, [])
        if free:
            return free.pop()
        return torch.empty(_native.get().mlp_workspace_bytes(self._shape, batch), dtype=torch.uint8, device=self._arena.device)

    def _give_workspace(self, batch: int, ws):
        free = self._ws_pool.setdefault(batch, [])
        if len(free) < 2:
            free.append(ws)

    def _check_arena(self):
        if self._arena is None:
            raise RuntimeError("Model runs only on a CUDA device (sm_100a kernels, no CPU fallback): call .cuda() first")
        base = self._arena.data_ptr()
        for p, (o, _) in zip(self._param_list, self._layout):
            if p.data_ptr() != base + 4 * o:
                self._flatten()
                break

    def _get_workspace(self, batch: int):
        if batch % 4:
            raise RuntimeError(f"Model needs a batch that is a multiple of 4, got {batch}")
        if self._workspace is None or self._workspace_batch != batch:
            nat = _native.get()
            self._workspace = torch.empty(nat.mlp_workspace_bytes(self._shape, batch), dtype=torch.uint8,
                                          device=self._arena.device)
            self._workspace_batch = batch
        return self._workspace

    @property
    def flat_params(self) -> torch.Tensor:
        self._check_arena()
        return self._arena

    @property
    def flat_grads(self) -> torch.Tensor:
        self._check_arena()
        return self._grad_arena

    def forward(self, x):
        self._check_arena()
        if not x.is_cuda:
            raise RuntimeError("Model.forward needs a CUDA tensor (no CPU fallback)")
        x = x.reshape(x.size(0), -1).to(torch.float32).contiguous()          # ref :151 x.view(x.size(0), -1)
        if not torch.is_grad_enabled():
            batch, padded = x.shape[0], -(-x.shape[0] // 4) * 4
            logits = torch.empty(padded, 10, dtype=torch.float32, device=x.device)
            _native.get().mlp_forward(self._shape, self._arena, _pad_rows(x, padded), None, self._get_workspace(padded),
                                      out_logits=logits)
            return logits[:batch]
        return _MLPFunction.apply(x, self, *self._param_list)

    def loss_and_grads(self, x, target, loss_scale: float = 1.0):
        """Fused step body without autograd: mean cross-entropy of ``model(x)`` vs ``target`` (device scalar) and
        its gradients written into ``flat_grads`` (= every ``p.grad``).  Same arithmetic as
        ``criterion(model(x), target).backward()`` (ref :72-74)."""
        self._check_arena()
        nat = _native.get()
        x = x.reshape(x.size(0), -1).to(torch.float32).contiguous()
        ws = self._get_workspace(x.shape[0])
        loss = torch.empty(1, dtype=torch.float32, device=x.device)
        nat.mlp_forward(self._shape, self._arena, x, target, ws, out_loss=loss)
        nat.mlp_backward(self._shape, self._arena, None, target, loss_scale, x.shape[0], ws, self._grad_arena)
        for p, (o, n) in zip(self._param_list, self._layout):
            p.grad = self._grad_arena[o:o + n].view(p.shape)
        return loss
