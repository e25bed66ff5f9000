#!/bin/bash
# ncu --set full capture of the three MLP GEMM kernels (one launch each) + summary
cd "$(dirname "$0")/.."
timeout 120 python tools/mlp_once.py 4096 2 || exit 1
ncu --set full --clock-control none --import-source on -k regex:gemm3x -s 12 -c 12 -o gpurun_out/mlp_gemm_full -f python tools/mlp_once.py 4096 2 > gpurun_out/mlp_ncu_full.log 2>&1
python tools/ncu_summary.py gpurun_out/mlp_gemm_full.ncu-rep 2>&1 | tail -14
