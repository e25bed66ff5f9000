import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_examples_b200 import _native
nat = _native.get(); nat.bind_device(0)
B = 4096
arena = torch.randn(21848, device="cuda") * 0.05
x = torch.randn(B, 1, 28, 28, device="cuda"); y = torch.randint(0, 10, (B,), device="cuda")
ws = torch.zeros(nat.workspace_bytes(B), dtype=torch.uint8, device="cuda"); sc = torch.zeros(nat.scratch_bytes(B), dtype=torch.uint8, device="cuda")
logp = torch.zeros(B, 10, device="cuda"); loss = torch.zeros(1, device="cuda")
nat.prof_enable(True)
acc = {}
for i in range(13):
    nat.net_forward(arena, x, y, None, None, 1, i, True, ws, logp, out_loss=loss, scratch=sc)
    torch.cuda.synchronize()
    # forward-only profile: events 0..3 are recorded by forward
    import ctypes
    ev = (ctypes.c_float * 8)()
torch.cuda.synchronize()
e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
nat.prof_enable(False)
e[0].record()
for i in range(20):
    nat.net_forward(arena, x, y, None, None, 1, i, True, ws, logp, out_loss=loss, scratch=sc)
e[1].record(); torch.cuda.synchronize()
print(os.environ.get("DDP_TC_DEBUG", "0"), os.environ.get("DDP_CONV2_FFMA", "-"), "fwd total us:", e[0].elapsed_time(e[1]) / 20 * 1e3)
