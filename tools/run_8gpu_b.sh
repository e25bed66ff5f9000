#!/bin/bash
# 8-GPU validation of the peer-memory allreduce path + scaling numbers.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
run() { timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port "$2" "${@:3}"; }
run 8 29711 tests/multi_gpu_worker.py > gpurun_out/multi_n8b.log 2>&1; echo "multi_gpu_worker exit=$?"; grep -E "rank 0\]|MULTI_GPU_OK|Error" gpurun_out/multi_n8b.log | tail -8
for n in 8 4 2; do
  run $n $((29720+n)) bench.py --gpus $n --steps 1500 --warmup 20 --no-extras 2> gpurun_out/scaleb_n$n.err | grep '^{' > gpurun_out/scaleb_n$n.json; echo "bench N=$n exit=$?"; cut -c1-200 gpurun_out/scaleb_n$n.json
done
run 8 29731 bench.py --gpus 8 --steps 1500 --warmup 20 --batch 512 --no-extras 2> gpurun_out/scaleb_n8_b512.err | grep '^{' > gpurun_out/scaleb_n8_b512.json; cut -c1-200 gpurun_out/scaleb_n8_b512.json
timeout 200 python bench.py --steps 1500 --warmup 20 --no-extras 2> gpurun_out/scaleb_n1.err | grep '^{' > gpurun_out/scaleb_n1.json; cut -c1-200 gpurun_out/scaleb_n1.json
