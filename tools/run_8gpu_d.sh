#!/bin/bash
# 8-GPU end-of-round check: multi-GPU worker (elastic latency, native MLP under DDP, symmetric allreduce) + the two
# headline 8-GPU bench configurations (B=4096/GPU weak scaling, B=512/GPU = BASELINE configs[2]).
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
N=${NGPU:-8}
run() { timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port "$2" "${@:3}"; }
run $N 29711 tests/multi_gpu_worker.py > gpurun_out/multi_n${N}d.log 2>&1; echo "multi_gpu_worker exit=$?"; grep -E "rank 0\]|MULTI_GPU_OK|Error" gpurun_out/multi_n${N}d.log | tail -10
run $N 29721 bench.py --gpus $N --steps 1500 --warmup 20 --no-extras 2> gpurun_out/scaled_n$N.err | grep '^{' > gpurun_out/scaled_n$N.json; echo "bench N=$N exit=$?"; cut -c1-220 gpurun_out/scaled_n$N.json
run $N 29731 bench.py --gpus $N --steps 1500 --warmup 20 --batch 512 --no-extras 2> gpurun_out/scaled_n${N}_b512.err | grep '^{' > gpurun_out/scaled_n${N}_b512.json; cut -c1-220 gpurun_out/scaled_n${N}_b512.json
