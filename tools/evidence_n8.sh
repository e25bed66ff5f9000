#!/bin/bash
# Multi-GPU end-of-round evidence (run under `gpurun --gpus 8`): the multi-rank worker (DDP hooks, peer / in-switch
# allreduce, elastic resize, lost-peer recovery) and the bench lines at N = 8 (B = 4096 with extras, B = 512), 4, 2, 1.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 400 $TR --nproc-per-node=8 --master-port 29541 tests/multi_gpu_worker.py > gpurun_out/multi_n8.log 2>&1; echo rc=$? >> gpurun_out/multi_n8.log
grep -E "rank 0|rc=|Error|assert|MULTI" gpurun_out/multi_n8.log | tail -14
timeout 400 $TR --nproc-per-node=8 --master-port 29542 bench.py --gpus 8 --steps 1000 --warmup 50 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo n8 rc=$?
timeout 150 $TR --nproc-per-node=8 --master-port 29543 bench.py --gpus 8 --steps 1000 --warmup 50 --batch 512 --no-extras > gpurun_out/bench_n8_b512.json 2> gpurun_out/bench_n8_b512.err
timeout 150 $TR --nproc-per-node=4 --master-port 29544 bench.py --gpus 4 --steps 1000 --warmup 50 --no-extras > gpurun_out/bench_n4.json 2> gpurun_out/bench_n4.err
timeout 150 $TR --nproc-per-node=2 --master-port 29545 bench.py --gpus 2 --steps 1000 --warmup 50 --no-extras > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
timeout 100 python bench.py --steps 1000 --warmup 50 --no-extras > gpurun_out/bench_n1b.json 2> gpurun_out/bench_n1b.err
for f in bench_n1b bench_n2 bench_n4 bench_n8 bench_n8_b512; do python tools/show_bench.py gpurun_out/$f.json | head -1 | cut -c1-120; done
