#!/bin/bash
# Two GPUs, after the round's last single-GPU kernel changes: the N-rank gradient check of both exchanges on hardware
# and the default weak-scaling bench line (render leg skipped to keep the visit short).
TAG=${1:-r02_final2}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $RUN tests/multi_gpu_check.py > gpurun_out/${TAG}_multi_gpu_check_n2.log 2>&1
echo "multi_gpu_check exit $?: $(tail -1 gpurun_out/${TAG}_multi_gpu_check_n2.log)"
timeout 400 $RUN bench.py --gpus 2 --steps 20 --warmup 5 --no-render --no-knn --no-cpu-baseline --no-drop-in > gpurun_out/${TAG}_bench_n2.json 2> gpurun_out/${TAG}_bench_n2.err
echo "bench N=2 exit $?"
python - <<PY
import json
try:
    j = json.load(open("gpurun_out/${TAG}_bench_n2.json"))
    print("value %.1f ms/step %.3f e2e %s exch %s clocks %s" % (j["value"], j["ms_per_step"], (j.get("e2e") or {}).get("value"), j.get("exchange_ms_rank0"), j.get("clocks")))
except Exception as e:
    print("FAILED", e, open("gpurun_out/${TAG}_bench_n2.err").read()[-1500:])
PY
