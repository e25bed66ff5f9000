"""Development probe for the symmetric two-shot allreduce: per-phase NVSwitch bandwidth (tools/probe/mc_probe.cu)
and NCCL / product-kernel time as a function of message size.
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -shared -Xcompiler -fPIC -o tools/probe/libmc_probe.so tools/probe/mc_probe.cu
    torchrun --nproc-per-node=N tools/sym_probe.py"""
import ctypes
import json
import os
import statistics
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import pytorch_distributed_examples_b200 as pkg
    hvd = pkg.hvd
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    hvd.init()
    c = hvd._require()
    N, R = c.world, c.rank
    lib = ctypes.CDLL(os.path.join(ROOT, "tools", "probe", "libmc_probe.so"))
    lib.mc_probe_launch.argtypes = [ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64), ctypes.c_int, ctypes.c_int, ctypes.c_int64,
                                    ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    big = 512 << 20
    assert c.enable_symmetric(big), c.sym_error
    sink = torch.zeros(4, device="cuda")
    reps_arr = (ctypes.c_uint64 * N)(*c.sym["replicas"])

    def timed(fn, reps=8, warm=2):
        ts = []
        for i in range(reps + warm):
            dist.barrier(); torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b)], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if i >= warm:
                ts.append(t.item())
        return statistics.median(ts) * 1e-3

    out = {"n_gpus": N}
    # 1. size sweep: NCCL vs product kernel
    for mb in (16, 102, 512):
        n = mb * (1 << 20) // 4 // 8 * 8
        S = n * 4
        t = torch.zeros(n, device="cuda")
        t_nccl = timed(lambda: c.allreduce_sum_(t))
        row = {"nccl_us": t_nccl * 1e6, "nccl_busbw": S / t_nccl * 2 * (N - 1) / N / 1e9}
        for ctas, un in ((64, 2), (64, 4), (32, 4), (148, 2)):
            ts = timed(lambda: c.sym_allreduce_(torch.float32, n, 0, 1.0 / N, max_ctas=ctas, unroll=un))
            row[f"sym_mc_ctas{ctas}_unroll{un}"] = {"us": ts * 1e6, "busbw": S / ts * 2 * (N - 1) / N / 1e9}
        out[f"{mb}MB"] = row
        del t
    # 2. phases at 102 MB (no cross-rank sync inside the kernel; chunk per rank)
    nvec = 102 * (1 << 20) // 16
    S = nvec * 16
    stream = torch.cuda.current_stream().cuda_stream
    names = {0: "ld_reduce+st", 1: "ld_reduce only", 2: "multimem.st only", 3: "ld_reduce+st slab/CTA", 4: "peer ld only", 5: "peer st only"}
    ph = {}
    for mode in (0, 1, 2, 3, 4, 5):
        for ctas, un in ((64, 4), (148, 4), (32, 8), (64, 2)):
            t = timed(lambda: lib.mc_probe_launch(c.sym["multicast"], reps_arr, N, R, nvec, mode, ctas, un, 0, sink.data_ptr(), stream))
            ph[f"{names[mode]} ctas{ctas} unroll{un}"] = {"us": t * 1e6, "chunk_GBps": S / N / t / 1e9}
    out["phases_102MB"] = ph
    if R == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
