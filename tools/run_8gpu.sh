#!/bin/bash
# One 8-GPU session: elastic 8->4->8 + hooks + in-graph allreduce (tests/multi_gpu_worker.py), the ResNet-50-size
# bucket path (tools/bucket_bench.py), and the scaling bench.  Outputs under gpurun_out/.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}
run() { timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port "$2" "${@:3}"; }
run $N 29701 tests/multi_gpu_worker.py > gpurun_out/multi_n$N.log 2>&1; echo "multi_gpu_worker exit=$?"; grep -E "rank 0\]|MULTI_GPU_OK|Error" gpurun_out/multi_n$N.log | tail -8
run $N 29702 tools/bucket_bench.py > gpurun_out/bucket_n$N.json 2> gpurun_out/bucket_n$N.err; echo "bucket_bench exit=$?"; tail -c 1500 gpurun_out/bucket_n$N.json
run $N 29703 bench.py --gpus $N --steps 1000 --warmup 20 --no-extras > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; echo "bench N=$N exit=$?"; cut -c1-330 gpurun_out/scale_n$N.json
run $N 29704 bench.py --gpus $N --steps 1000 --warmup 20 --batch 512 --no-extras > gpurun_out/scale_n${N}_b512.json 2> gpurun_out/scale_n${N}_b512.err; echo "bench N=$N B=512 exit=$?"; cut -c1-330 gpurun_out/scale_n${N}_b512.json
if [ "$N" -ge 8 ]; then
  run 4 29705 bench.py --gpus 4 --steps 1000 --warmup 20 --no-extras > gpurun_out/scale_n4.json 2> gpurun_out/scale_n4.err; echo "bench N=4 exit=$?"; cut -c1-330 gpurun_out/scale_n4.json
fi
