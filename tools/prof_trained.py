"""Program for ncu / sparsity statistics on a TRAINED state: TrainStep for STEPS steps on the synthetic digits
(as bench.py does), then REPS eager fwd+bwd passes through the C-ABI.  Prints the live fractions the sparse
backward kernels depend on.   python tools/prof_trained.py [B=4096] [STEPS=300] [REPS=3]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pytorch_distributed_examples_b200 as pkg  # noqa: E402
from pytorch_distributed_examples_b200 import _native  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
STEPS = int(sys.argv[2]) if len(sys.argv) > 2 else 300
REPS = int(sys.argv[3]) if len(sys.argv) > 3 else 3
torch.cuda.set_device(0)
nat = _native.get()
torch.manual_seed(0)
model = pkg.Net().cuda().train()
opt = pkg.FusedOptimizer(torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9))
slots = 4
step = pkg.TrainStep(model, opt, B, slots=slots)
ds = pkg.SyntheticMNIST(length=1 << 30, seed=1234)
for s in range(slots):
    nat.synth_batch(ds.seed, None, s * B, B, step.xs[s], step.ys[s])
for i in range(STEPS):
    step.run_slot(i % slots)
torch.cuda.synchronize()
print("loss after", STEPS, "steps:", float(step.loss.item()))
for i in range(REPS):
    nat.net_forward(model._arena, step.xs[0], step.ys[0], None, None, 1, i, True, step.ws, step.logp,
                    out_loss=step.loss, scratch=step.scratch)
    nat.net_backward(model._arena, step.xs[0], step.logp, None, step.ys[0], 1.0, True, step.ws, step.scratch,
                     model._grad_arena)
torch.cuda.synchronize()


def a256(v):
    return (v + 255) & ~255


ws = step.ws
o_a1 = 0
o_idx1 = a256(B * 1440 * 4)
o_a2 = o_idx1 + a256(B * 1440)
a1 = ws[o_a1:o_a1 + B * 1440 * 4].view(torch.float32)
a2 = ws[o_a2:o_a2 + B * 320 * 4].view(torch.float32).view(B, 20, 16)
print("a1 > 0 fraction:", float((a1 > 0).float().mean()))
print("a2 > 0 fraction:", float((a2 > 0).float().mean()), " live entries/sample:", float((a2 > 0).float().sum() / B))
print("channels with any live entry per sample:", float((a2 > 0).any(dim=2).float().sum() / B))
