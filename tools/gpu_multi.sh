#!/bin/bash
# Multi-GPU visit: hardware correctness of both exchanges, then bench at N ranks with the peer exchange and with NCCL.
# Usage: bash tools/gpu_multi.sh <tag> <N> [extra bench args]
TAG=$1; N=$2; shift; shift
mkdir -p gpurun_out
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 "$@"; }
run tests/multi_gpu_check.py > gpurun_out/${TAG}_check.log 2>&1; echo "check exit $?"; grep -h "MULTI_GPU_CHECK\|MISMATCH\|Error\|error" gpurun_out/${TAG}_check.log | head -10
run tests/multi_gpu_check.py --views $((2*N)) > gpurun_out/${TAG}_check2.log 2>&1; echo "check2 exit $?"; grep -h "MULTI_GPU_CHECK\|MISMATCH\|Error\|error" gpurun_out/${TAG}_check2.log | head -10
for mode in peer nccl; do
  GSM_VP_EXCHANGE=$mode run bench.py --gpus $N --steps 20 --warmup 5 --no-render "$@" > gpurun_out/${TAG}_bench_n${N}_${mode}.json 2> gpurun_out/${TAG}_bench_n${N}_${mode}.err
  echo "bench $mode exit $?"
  python - <<PY
import json
try:
    j = json.load(open("gpurun_out/${TAG}_bench_n${N}_${mode}.json"))
    print("$mode N=$N value %.1f ms %.3f e2e %s" % (j["value"], j["ms_per_step"], (j.get("e2e") or {}).get("ms_per_step")), j["config"]["exchange"][:60])
    print("   ", " ".join("%s=%.3f" % (r["kernel"], r["ms"]) for r in j["stages"]))
except Exception as e:
    print("$mode FAILED", e, open("gpurun_out/${TAG}_bench_n${N}_${mode}.err").read()[-1500:])
PY
done
