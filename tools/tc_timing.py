"""Per-role clock64 counters of conv2_fwd_tc (setup / producer / MMA issuer / epilogue, waits and work) on one GPU.
Needs the instrumented build of the library next to the shipped one:
    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -DDDP_TC_TIMING -shared -Xcompiler -fPIC \
         -I<nccl include> pytorch_distributed_examples_b200/csrc/*.cu -o pytorch_distributed_examples_b200/lib/libddp_step_dbg.so -ldl
The counters cost a global read-modify-write each, so the instrumented kernel's total is not a benchmark number."""
import ctypes, os, sys, torch
os.environ["DDP_LIB_PATH"] = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "pytorch_distributed_examples_b200/lib/libddp_step_dbg.so")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_examples_b200 import _native
nat = _native.get(); nat.bind_device(0)
B = 4096
arena = torch.randn(21848, device="cuda") * 0.05
x = torch.randn(B, 1, 28, 28, device="cuda"); y = torch.randint(0, 10, (B,), device="cuda")
ws = torch.zeros(nat.workspace_bytes(B), dtype=torch.uint8, device="cuda"); sc = torch.zeros(nat.scratch_bytes(B), dtype=torch.uint8, device="cuda")
logp = torch.zeros(B, 10, device="cuda"); loss = torch.zeros(1, device="cuda")
for i in range(3):
    nat.net_forward(arena, x, y, None, None, 1, i, True, ws, logp, out_loss=loss, scratch=sc)
torch.cuda.synchronize()
nat.lib.ddp_debug_tc_timing.argtypes = [ctypes.c_void_p, ctypes.c_int]
nat.lib.ddp_debug_tc_timing(None, 1)
nat.net_forward(arena, x, y, None, None, 1, 9, True, ws, logp, out_loss=loss, scratch=sc)
torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * (160 * 16))()
nat.lib.ddp_debug_tc_timing(buf, 0)
names = ["setup", "prod wait raw", "prod wait empty", "prod build", "prod fence+arrive", "mma wait tmem", "mma wait full", "mma issue", "epi wait acc", "epi work", "kernel total"]
for cta in (0, 1, 73, 147):
    print("CTA", cta, {n: buf[cta * 16 + i] for i, n in enumerate(names)})
