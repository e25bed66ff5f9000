"""Minimal program for ncu: a few eager fwd+bwd passes of the CNN at B (default 4096) through the C-ABI."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pytorch_distributed_examples_b200 import _native  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
REPS = int(sys.argv[2]) if len(sys.argv) > 2 else 4
nat = _native.get()
nat.bind_device(0)
torch.manual_seed(0)
arena = torch.randn(21848, device="cuda") * 0.05
x = torch.randn(B, 1, 28, 28, device="cuda")
y = torch.randint(0, 10, (B,), device="cuda")
ws = torch.zeros(nat.workspace_bytes(B), dtype=torch.uint8, device="cuda")
sc = torch.zeros(nat.scratch_bytes(B), dtype=torch.uint8, device="cuda")
logp = torch.zeros(B, 10, device="cuda"); loss = torch.zeros(1, device="cuda"); bucket = torch.zeros(21848, device="cuda")
for i in range(REPS):
    nat.net_forward(arena, x, y, None, None, 1, i, True, ws, logp, out_loss=loss, scratch=sc)
    nat.net_backward(arena, x, logp, None, y, 1.0, True, ws, sc, bucket)
torch.cuda.synchronize()
print("loss", loss.item())
