"""Debug aid: one forward / backward at batch B through the C-ABI with a print after every host call."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pytorch_distributed_examples_b200 import _native
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
nat = _native.get(); nat.bind_device(0)
torch.manual_seed(0)
arena = torch.randn(21848, device="cuda") * 0.05
x = torch.randn(B, 1, 28, 28, device="cuda"); y = torch.randint(0, 10, (B,), device="cuda")
ws = torch.zeros(nat.workspace_bytes(B), dtype=torch.uint8, device="cuda")
sc = torch.zeros(nat.scratch_bytes(B), dtype=torch.uint8, device="cuda")
logp = torch.zeros(B, 10, device="cuda"); loss = torch.zeros(1, device="cuda"); bucket = torch.zeros(21848, device="cuda")
torch.cuda.synchronize(); print("setup ok", flush=True)
nat.net_forward(arena, x, y, None, None, 1, 0, True, ws, logp, out_loss=loss, scratch=sc); print("fwd enqueued", flush=True)
torch.cuda.synchronize(); print("fwd done", loss.item(), flush=True)
nat.net_backward(arena, x, logp, None, y, 1.0, True, ws, sc, bucket); print("bwd enqueued", flush=True)
torch.cuda.synchronize(); print("bwd done", bucket.abs().sum().item(), flush=True)
if hasattr(nat, "net_train_fwdbwd"):
    nat.net_train_fwdbwd(arena, x, y, None, None, 1, 0, ws, logp, loss, sc, bucket); print("fused enqueued", flush=True)
    torch.cuda.synchronize(); print("fused done", loss.item(), bucket.abs().sum().item(), flush=True)
