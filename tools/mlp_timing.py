"""Per-role clock64 breakdown of the MLP GEMM kernel.  Needs a debug build of the library next to the real one:
    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC -DDDP_MLP_TIMING \
         -I<site-packages>/nvidia/nccl/include pytorch_distributed_examples_b200/csrc/*.cu \
         -o pytorch_distributed_examples_b200/lib/libddp_step_dbg.so -ldl
(the counters live in registers and are flushed once per CTA, so the instrumented kernel runs at full speed)."""
import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["DDP_LIB_PATH"] = os.path.join(ROOT, "pytorch_distributed_examples_b200/lib/libddp_step_dbg%s.so" % (sys.argv[1] if len(sys.argv) > 1 else ""))
sys.path.insert(0, ROOT)
from pytorch_distributed_examples_b200 import _native
nat = _native.get(); nat.bind_device(0)
nat.lib.ddp_debug_mlp_timing.argtypes = [ctypes.c_void_p, ctypes.c_int]
names = ["kernel", "prod wait empty", "mma wait tmem", "mma wait full", "mma issue", "epi wait acc", "epi drain", "epi store", "t first operands", "t last mma issued"]
def run(tag, fn):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    nat.lib.ddp_debug_mlp_timing(None, 1)
    fn(); torch.cuda.synchronize()
    buf = (ctypes.c_ulonglong * (160 * 16))()
    nat.lib.ddp_debug_mlp_timing(buf, 0)
    for cta in (0, 73, 147):
        print(tag, "CTA", cta, {n: buf[cta * 16 + i] for i, n in enumerate(names)})
M, N, K = 4096, 1024, 1024
a = torch.randn(M, K, device="cuda"); b = torch.randn(N, K, device="cuda")
ah, al = nat.split_tf32(a); bh, bl = nat.split_tf32(b)
out = torch.empty(M, N, device="cuda")
run("plain 4096x1024x1024", lambda: nat.gemm_3xtf32(ah, al, bh, bl, out=out))
a = torch.randn(1024, 4096, device="cuda"); b = torch.randn(1024, 4096, device="cuda")
ah, al = nat.split_tf32(a); bh, bl = nat.split_tf32(b)
out2 = torch.empty(1024, 1024, device="cuda")
run("plain 1024x1024x4096", lambda: nat.gemm_3xtf32(ah, al, bh, bl, out=out2))
# the model's forward GEMMs (bias+relu+split epilogue): time one whole forward
import pytorch_distributed_examples_b200 as pkg
model = pkg.Model(5, 1024).cuda()
x = torch.randn(4096, 784, device="cuda"); y = torch.randint(0, 10, (4096,), device="cuda")
with torch.no_grad():
    run("forward (6 GEMMs, bias+relu+split)", lambda: model(x))
run("forward+backward (17 GEMM launches)", lambda: model.loss_and_grads(x, y))
