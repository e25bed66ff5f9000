#!/bin/bash
# 8-GPU validation + bandwidth of the symmetric-memory two-shot allreduce (csrc/sym_allreduce.cu).
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
run() { timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port "$2" "${@:3}"; }
run ${NGPU:-8} 29711 tests/multi_gpu_worker.py > gpurun_out/multi_n8c.log 2>&1; echo "multi_gpu_worker exit=$?"; grep -E "rank 0\]|MULTI_GPU_OK|Error" gpurun_out/multi_n8c.log | tail -8
run ${NGPU:-8} 29712 tools/bucket_bench.py 2> gpurun_out/bucket_n8c.err | grep '^{' > gpurun_out/bucket_n8c.json; echo "bucket_bench exit=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/bucket_n8c.json"))
for k in ("fp32", "bf16"):
    print(k, "nccl single-shot us", round(d[k]["allreduce_us"], 1), "busbw", round(d[k]["allreduce_busbw_GBps"], 1), "whole", round(d[k]["flatten+allreduce+update_us"], 1))
    for m, v in d[k]["symmetric_two_shot"].items():
        if isinstance(v, dict) and "single_shot" in v:
            print("  ", m, "single", round(v["single_shot"]["us"], 1), round(v["single_shot"]["busbw_GBps"], 1), "| b2b", round(v["back_to_back"]["us"], 1), round(v["back_to_back"]["busbw_GBps"], 1))
        elif isinstance(v, dict):
            print("  ", m, round(v["us"], 1), round(v["busbw_GBps"], 1))
        else:
            print("  ", m, v)
PY
N=${NGPU:-8}
if [ "${RESNET:-1}" = "1" ]; then
  run $N 29714 tools/resnet_ddp_bench.py 2> gpurun_out/resnet_n${N}c.err | grep '^{' > gpurun_out/resnet_n${N}c.json; echo "resnet exit=$?"
  python -c "
import json; d=json.load(open('gpurun_out/resnet_n${N}c.json'))
for k,v in d.items(): print(k, v)"
fi
