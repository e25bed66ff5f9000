#!/bin/bash
# One-GPU end-of-round evidence (run under gpurun): full GPU test suite, the default bench line, the ncu launch list of
# the bench command and one `ncu --set full` capture of every kernel of the step on a trained state.
# Each ncu pass runs only after the same command has exited 0 without ncu; numbers printed under ncu are never bench values.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -u -m pytest tests -m gpu -x -q -o faulthandler_timeout=200 2>&1 | tail -4 > gpurun_out/gputest.log; cat gpurun_out/gputest.log
timeout 400 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 120 python bench.py --steps 20 --warmup 5 --no-extras > gpurun_out/plain1.log 2>&1 && \
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 20 --warmup 5 --no-extras > gpurun_out/ncu1.log 2>&1
python tools/prof_trained.py 4096 300 3 > gpurun_out/plain2.log 2>&1 && \
  timeout 400 ncu --set full --clock-control none --import-source on --graph-profiling graph -s 27 -c 9 -o gpurun_out/prof_r02 \
    python tools/prof_trained.py 4096 300 3 > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
